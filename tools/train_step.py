"""One training step of the drop-in ArtSpeech (forward with saved activations, loss, backward, SGD step) at a
training-sized batch: the target of the ncu launch list that shows which kernels a training step runs."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import json
import time
import torch
from artspeech_b200 import _lib, models, synthetic

B, T = 32, 200
torch.manual_seed(0)
net = models.ArtSpeech(45, 10, dropout=0.2, precision="fp32").cuda().train()
opt = torch.optim.SGD(net.parameters(), lr=1e-3)
x = synthetic.tokens(B, T, 45, seed=1).cuda()
L = [int(v) for v in synthetic.sorted_lengths(B, 80, T, seed=1)]
tgt = synthetic.contours(B * T, 10, 50, seed=2).view(B, T, 10, 2, 50).cuda()
times = []
with torch.backends.cudnn.flags(enabled=False):
    for it in range(4):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        l0 = _lib.launch_count()
        opt.zero_grad()
        out = net(x, L)
        loss = (out - tgt).square().mean()
        loss.backward()
        opt.step()
        torch.cuda.synchronize()
        times.append(time.perf_counter() - t0)
        launches = _lib.launch_count() - l0
print(json.dumps({"batch": B, "seq_len": T, "ms_per_training_step": min(times) * 1e3, "library_launches_per_step": launches,
                  "frames_per_s": B * T / min(times), "loss": float(loss)}))
