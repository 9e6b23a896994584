"""Does the forked tract-variable branch of the captured evaluation step overlap the forward pass?
Times the CUDA-graph replay of evaluate_batch with and without the side stream (L2 flushed between
replays, CUDA events)."""
import json
import sys

import torch

sys.path.insert(0, ".")
from artspeech_b200 import evaluation, models, synthetic  # noqa: E402


def main():
    dev = torch.device("cuda", 0)
    B, T = 256, 100
    torch.manual_seed(0)
    net = models.ArtSpeech(45, 10).to(dev).eval()
    tok = synthetic.tokens(B, T, 45, seed=100).to(dev)
    tgt = synthetic.contours(B * T, 10, 50, seed=200).view(B, T, 10, 2, 50).to(dev)
    uinc = synthetic.UPPER_INCISOR_REF.to(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    res = {}
    for name, ov in (("serial", False), ("forked", True), ("serial2", False), ("forked2", True)):
        g = evaluation.GraphedEvaluator(net, tok, [T] * B, tgt, upper_incisor=uinc, overlap_target=ov)
        for _ in range(5):
            g()
        ts = []
        for _ in range(30):
            flush.zero_()
            s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
            s.record(); g(); e.record()
            torch.cuda.synchronize()
            ts.append(s.elapsed_time(e))
        ts.sort()
        res[name] = {"median_ms": ts[len(ts) // 2], "min_ms": ts[0]}
    print(json.dumps(res))


if __name__ == "__main__":
    main()
