"""One small pass over every kernel of the library (target of compute-sanitizer runs): evaluation step
(forward, P2CP, tract variables, accumulation) for both cells and both precisions, the standalone predictor,
SimpleArtSpeech, the principal-components model, the area function, and one training step (forward with
saved activations + backward, dropout on)."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from artspeech_b200 import evaluation, models, synthetic
from artspeech_b200 import principal_components as pc
import artspeech_b200.area_function as AF

for cell in ("gru", "lstm"):
    for precision in ("fast", "fp32"):
        net = models.ArtSpeech(45, 10, precision=precision, rnn=cell).cuda().eval()
        for (B, T) in [(5, 23), (37, 61), (300, 7)]:
            x = synthetic.tokens(B, T, 45, seed=B)
            L = synthetic.sorted_lengths(B, max(1, T // 4), T, seed=B)
            tgt = synthetic.contours(B * T, 10, 50, seed=B).view(B, T, 10, 2, 50)
            with torch.no_grad():
                res = evaluation.evaluate_batch(net, x.cuda(), L, tgt.cuda())
            torch.cuda.synchronize()
            assert not torch.isnan(res["outputs"]).any()
            evaluation.sentence_metrics_device(res["outputs"], tgt.cuda(), torch.as_tensor(L, dtype=torch.int32).cuda())
pred = models.ArticulatorPredictor(64, 50).cuda()
with torch.no_grad():
    pred(torch.randn(3, 70, 64, device="cuda"))
simple = models.SimpleArtSpeech(45, 10).cuda().eval()
with torch.no_grad():
    simple(synthetic.tokens(4, 9, 45).cuda(), None)
pcm = pc.PrincipalComponentsArtSpeech(45, {"a": 4, "b": 5}, rnn="LSTM").cuda().eval()
pcm(synthetic.tokens(5, 23, 45, seed=1).cuda(), [23, 20, 11, 5, 2])
wi, we = synthetic.walls(300, seed=2)
grid = AF.DeviceGrid(AF.build_semipolar_grid(**synthetic.AREA_GRID_ARGS), torch.device("cuda"))
AF.area_function_batch(wi.cuda(), we.cuda(), grid)
# one training step, dropout on
train = models.ArtSpeech(45, 3, dropout=0.3, precision="fp32").cuda().train()
out = train(synthetic.tokens(11, 17, 45, seed=4).cuda(), [17, 17, 16, 12, 12, 9, 5, 3, 2, 1, 1])
out.square().mean().backward()
torch.cuda.synchronize()
print("ok")
