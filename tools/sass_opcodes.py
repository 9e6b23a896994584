"""SASS opcode counts per kernel: cuobjdump -sass artspeech_b200/libartspeech_b200.so | c++filt | python tools/sass_opcodes.py"""
import collections
import re
import sys

KEYS = ["UTCxMMA", "LDTM", "STTM", "UBLKCP", "UTMALDG", "LDGSTS", "packed_fp32x2", "FMNMX3", "HMMA"]
cur, cnt = None, collections.OrderedDict()
for line in sys.stdin:
    m = re.search(r"Function : (.*)", line)
    if m:
        name = re.sub(r"\(anonymous namespace\)::", "", m.group(1).strip())
        cur = re.sub(r"\(.*", "", name)
        cnt.setdefault(cur, collections.Counter())
        continue
    if cur is None:
        continue
    m = re.search(r"^\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_.]+)", line)
    if not m:
        continue
    op, c = m.group(1), cnt[cur]
    if op.startswith("UTC") and "MMA" in op:
        c["UTCxMMA"] += 1
    for k in ("LDTM", "STTM", "UBLKCP", "UTMALDG", "LDGSTS", "FMNMX3", "HMMA"):
        if op.startswith(k):
            c[k] += 1
    if op.split(".")[0] in ("FFMA2", "FADD2", "FMUL2"):
        c["packed_fp32x2"] += 1
print("# SASS opcode counts per kernel in artspeech_b200/libartspeech_b200.so (cuobjdump -sass | c++filt | tools/sass_opcodes.py).")
print("# UTCxMMA = tcgen05.mma (UTCHMMA/UTCQMMA...), LDTM/STTM = tcgen05.ld/st, UBLKCP = cp.async.bulk (1-D TMA), UTMALDG = tensor-map TMA")
print("# (none: every staged tile here is a contiguous 1-D range), LDGSTS = cp.async, HMMA = mma.sync (the fp32-precision 3xTF32 kernels).")
for k, c in sorted(cnt.items()):
    if sum(c.values()):
        print(k, " ".join(f"{q}={c[q]}" for q in KEYS))
