"""A few forward passes at configs[1] (batch 256 x 100) - the target of single-kernel ncu captures."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from artspeech_b200 import models, synthetic

cell = sys.argv[1] if len(sys.argv) > 1 else "gru"
net = models.ArtSpeech(45, 10, precision="fast", rnn=cell).cuda().eval()
x = synthetic.tokens(256, 100, 45, seed=1).cuda()
L = [100] * 256
with torch.no_grad():
    for _ in range(3):
        out = net(x, L)
torch.cuda.synchronize()
print("ok", float(out.mean()))
