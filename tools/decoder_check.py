"""Compare the tcgen05 decoder with the mma.sync decoder and the CPU oracle (small + config-2 sizes)."""
import sys
import torch
sys.path.insert(0, ".")
from artspeech_b200 import models, synthetic
from oracle import model as omodel

sd = omodel.randomize_affine(omodel.make_state_dict(45, 10, seed=0), seed=3)
net = models.ArtSpeech(45, 10, precision="fast")
net.load_state_dict(sd)
net = net.cuda().eval()
for (B, T) in [(3, 12), (37, 61), (256, 100)]:
    x = synthetic.tokens(B, T, 45, seed=5)
    L = synthetic.sorted_lengths(B, max(1, T // 3), T, seed=5)
    with torch.no_grad():
        net.set_decoder("mma")
        a = net(x.cuda(), L)
        net.set_decoder("tcgen05")
        b = net(x.cuda(), L)
        torch.cuda.synchronize()
    nb = min(B, 16)
    want = omodel.artspeech_forward(sd, x[:nb], L[:nb], torch.float32)
    Tw = want.shape[1]
    print(f"B={B} T={T}: tcgen05 vs mma {float((a - b).abs().max()):.3e}  "
          f"mma vs oracle {float((a[:nb, :Tw].cpu() - want).abs().max()):.3e}  "
          f"tcgen05 vs oracle {float((b[:nb, :Tw].cpu() - want).abs().max()):.3e}  nan={bool(torch.isnan(b).any())}")
