"""Three evaluation steps at configs[1] (batch 256 x 100: forward, P2CP, tract variables of prediction and
target, accumulation) - the target of `ncu --set full` captures of the step's kernels."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from artspeech_b200 import evaluation, models, synthetic

precision = sys.argv[1] if len(sys.argv) > 1 else "fast"
B, T = 256, 100
net = models.ArtSpeech(45, 10, precision=precision).cuda().eval()
x = synthetic.tokens(B, T, 45, seed=1).cuda()
L = [T] * B
tgt = synthetic.contours(B * T, 10, 50, seed=3, device="cuda").view(B, T, 10, 2, 50)
with torch.no_grad():
    for _ in range(3):
        res = evaluation.evaluate_batch(net, x, L, tgt, upper_incisor=synthetic.UPPER_INCISOR_REF)
torch.cuda.synchronize()
print("ok", float(res["acc"][0]))
