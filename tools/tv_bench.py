"""Tract-variable kernel alone: frames/s at F distinct frames (pred + target = two launches), CUDA events."""
import json
import sys

import torch

sys.path.insert(0, ".")
from artspeech_b200 import synthetic  # noqa: E402
from artspeech_b200.tract_variables import tract_variables_batch  # noqa: E402


def main():
    F = int(sys.argv[1]) if len(sys.argv) > 1 else 524288
    dev = torch.device("cuda")
    tgt = torch.cat([synthetic.contours(131072, 10, 50, seed=5000 + k, device=dev) for k in range(F // 131072)])
    pred = synthetic.perturbed(tgt, seed=7)
    uinc = synthetic.UPPER_INCISOR_REF.to(dev)
    fn = lambda: (tract_variables_batch(pred, synthetic.ARTICULATORS, uinc),
                  tract_variables_batch(tgt, synthetic.ARTICULATORS, uinc))
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
    ev[0].record()
    for i in range(5):
        fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    ts = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(5))
    ms = ts[2]
    print(json.dumps({"frames": F, "ms": ms, "frames_per_s": F / ms * 1e3, "gbs": F * 5024 / ms * 1e-6,
                      "clk_per_set_per_sm": ms * 1e-3 * 1.9e9 / (2 * F / 148)}))


if __name__ == "__main__":
    main()
