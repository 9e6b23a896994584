"""Per-stage time of the forward pass with the LSTM cell next to the GRU (configs[1] and configs[3] shapes)."""
import json
import sys
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from artspeech_b200 import models, synthetic

out = {}
for cell in ("gru", "lstm"):
    for (B, T) in [(256, 100), (64, 1024)]:
        net = models.ArtSpeech(45, 10, precision="fast", rnn=cell).cuda().eval()
        x = synthetic.tokens(B, T, 45, seed=1).cuda()
        L = [T] * B
        ld = torch.tensor(L, dtype=torch.int32, device="cuda")
        with torch.no_grad():
            for _ in range(5):
                net(x, L, lengths_dev=ld)
            net.profile(True)
            for _ in range(20):
                net(x, L, lengths_dev=ld)
            torch.cuda.synchronize()
            prof = net.read_profile()
        out[f"{cell}_{B}x{T}"] = {k: round(v[0] / max(1, v[1]), 4) for k, v in prof.items()}
        out[f"{cell}_{B}x{T}"]["us_per_step_l0"] = round(1e3 * out[f"{cell}_{B}x{T}"]["gru_l0"] / T, 3)
print(json.dumps(out, indent=1))
