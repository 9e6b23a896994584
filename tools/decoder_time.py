"""Decoder / per-stage times at configs[1] plus the parity of the tcgen05 decoder against the oracle."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from artspeech_b200 import _lib
if os.environ.get("ASB_LIB"):   # a tools/build_variant.py library
    _lib.LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), os.environ["ASB_LIB"])
from artspeech_b200 import models, synthetic
from oracle import model as omodel

sd = omodel.randomize_affine(omodel.make_state_dict(45, 10, seed=0), seed=3)
net = models.ArtSpeech(45, 10, precision="fast")
net.load_state_dict(sd)
net = net.cuda().eval()
B, T = 256, 100
x = synthetic.tokens(B, T, 45, seed=5)
L = synthetic.sorted_lengths(B, 40, T, seed=5)
with torch.no_grad():
    out = net(x.cuda(), L)
    net.set_decoder("mma")
    ref = net(x.cuda(), L)
    net.set_decoder("tcgen05")
want = omodel.artspeech_forward(sd, x[:8], L[:8], torch.float32)
print(f"tcgen05 vs mma {float((out - ref).abs().max()):.3e}  tcgen05 vs oracle "
      f"{float((out[:8, :want.shape[1]].cpu() - want).abs().max()):.3e}")
xd = x.cuda()
Lf = [T] * B
ld = torch.tensor(Lf, dtype=torch.int32, device="cuda")
with torch.no_grad():
    for _ in range(5):
        net(xd, Lf, lengths_dev=ld)
    net.profile(True)
    for _ in range(30):
        net(xd, Lf, lengths_dev=ld)
    torch.cuda.synchronize()
print({k: round(v[0] / max(1, v[1]), 4) for k, v in net.read_profile().items()})
