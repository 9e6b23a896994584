"""Time the 50x50 P2CP kernel alone (CUDA events) at the bench size and at a large size.

    python tools/p2cp_time.py [frames ...]     # default 25600 (bench: 256 x 100) and 262144
"""
import json
import sys

import torch

sys.path.insert(0, ".")
from artspeech_b200 import synthetic  # noqa: E402
import artspeech_b200.metrics as M  # noqa: E402


def timeit(fn, iters=20, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(iters + 1)]
    ev[0].record()
    for i in range(iters):
        fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    ts = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(iters))
    return ts[len(ts) // 2]


def main():
    sizes = [int(a) for a in sys.argv[1:]] or [25600, 262144]
    base = synthetic.tv_frames(1024, seed=1).cuda()[:, :6]
    for F in sizes:
        reps = (F + 1023) // 1024
        tgt = base.repeat(reps, 1, 1, 1)[:F].contiguous()
        pred = (tgt + 0.01 * torch.randn_like(tgt)).contiguous()
        ref = None
        out = M.p2cp_distance(pred, tgt)
        u = pred[:64].double(); v = tgt[:64].double()
        d = (u.unsqueeze(-1) - v.unsqueeze(-2)).pow(2).sum(-3).sqrt()  # (64, 6, 50, 50)
        ref = 0.5 * (d.min(-1).values.mean(-1) + d.min(-2).values.mean(-1))
        err = ((out[:64].double() - ref).abs() / ref).max().item()
        ms = timeit(lambda: M.p2cp_distance(pred, tgt))
        pairs = F * 6 * 2500
        print(json.dumps({"frames": F, "items": F * 6, "ms": round(ms, 4), "pairs_per_s": pairs / ms * 1e3,
                          "max_rel_err_vs_fp64": err}))


if __name__ == "__main__":
    main()
