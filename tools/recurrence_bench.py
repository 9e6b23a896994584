"""BASELINE configs[3] (batch 64, length 1024): per-stage times of the forward, us per recurrent step."""
import json
import sys

import torch

sys.path.insert(0, ".")
from artspeech_b200 import models, synthetic  # noqa: E402


def main():
    B, T = 64, 1024
    torch.manual_seed(0)
    net = models.ArtSpeech(45, 10, precision="fast").cuda().eval()
    x = synthetic.tokens(B, T, 45, seed=1).cuda()
    L = [T] * B
    res = {"batch": B, "seq_len": T}
    for impl in ("mma", "tcgen05"):
        net.set_recurrent(impl)
        with torch.no_grad():
            for _ in range(3):
                net(x, L)
            net.profile(True)
            net.read_profile(reset=True)
            for _ in range(5):
                net(x, L)
            torch.cuda.synchronize()
            prof = net.read_profile(reset=True)
            net.profile(False)
        ms = {k: v[0] / v[1] for k, v in prof.items()}
        res[impl] = {"stage_ms": {k: round(v, 4) for k, v in ms.items()},
                     "us_per_step_layer0": ms["gru_l0"] * 1e3 / T, "us_per_step_layer1": ms["gru_l1"] * 1e3 / T,
                     "forward_ms": sum(ms.values()), "frames_per_s": B * T / (sum(ms.values()) * 1e-3)}
    print(json.dumps(res))


if __name__ == "__main__":
    main()
