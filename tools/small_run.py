"""One small pass over every kernel of the hot path (target of compute-sanitizer runs)."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from artspeech_b200 import evaluation, models, synthetic
import artspeech_b200.area_function as AF

for cell in ("gru", "lstm"):
    net = models.ArtSpeech(45, 10, precision="fast", rnn=cell).cuda().eval()
    for (B, T) in [(5, 23), (37, 61), (300, 7)]:
        x = synthetic.tokens(B, T, 45, seed=B)
        L = synthetic.sorted_lengths(B, max(1, T // 4), T, seed=B)
        tgt = synthetic.contours(B * T, 10, 50, seed=B).view(B, T, 10, 2, 50)
        with torch.no_grad():
            res = evaluation.evaluate_batch(net, x.cuda(), L, tgt.cuda())
        torch.cuda.synchronize()
        assert not torch.isnan(res["outputs"]).any()
net32 = models.ArtSpeech(45, 10, precision="fp32").cuda().eval()
with torch.no_grad():
    net32(synthetic.tokens(5, 23, 45, seed=1).cuda(), [23, 20, 11, 5, 2])
wi, we = synthetic.walls(300, seed=2)
grid = AF.DeviceGrid(AF.build_semipolar_grid(**synthetic.AREA_GRID_ARGS), torch.device("cuda"))
AF.area_function_batch(wi.cuda(), we.cuda(), grid)
torch.cuda.synchronize()
print("ok")
