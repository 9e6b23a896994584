"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel.
usage: python tools/launch_summary.py launches.csv [steps]"""
import csv
import sys
from collections import OrderedDict


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    hdr, data = rows[hi], rows[hi + 1:]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = OrderedDict()
    for r in data:
        if len(r) <= vi:
            continue
        name = r[ki].split("(")[0][-64:]
        v = float(r[vi].replace(",", ""))
        v = v / 1000 if r[ui] == "ns" else (v * 1000 if r[ui] == "ms" else v)
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    print(f"# {len(data)} launches, {tot:.1f} us total device time (cold-cache, serialised: compare SHARES)")
    print(f"# {'total_us':>10} {'count':>6} {'share':>7} {'us/launch':>10}  kernel")
    for n, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f"  {t:10.1f} {c:6d} {100 * t / tot:6.1f}% {t / c:10.1f}  {n}")


if __name__ == "__main__":
    main()
