"""Device-resident step time of PipelinedEvaluator by number of steps in flight (configs[1])."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from artspeech_b200 import _lib
if os.environ.get("ASB_LIB"):   # a tools/build_variant.py library
    _lib.LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), os.environ["ASB_LIB"])
from artspeech_b200 import evaluation, models, synthetic

B, T = 256, 100
net = models.ArtSpeech(45, 10, precision=sys.argv[1] if len(sys.argv) > 1 else "fast").cuda().eval()
L = [T] * B
ld = torch.tensor(L, dtype=torch.int32, device="cuda")
sets = [(synthetic.tokens(B, T, 45, seed=s).cuda(),
         synthetic.contours(B * T, 10, 50, seed=10 + s, device="cuda").view(B, T, 10, 2, 50)) for s in range(2)]
for depth in ((3,) if os.environ.get("ASB_LIB") or len(sys.argv) > 2 else (1, 2, 3, 4)):
    pipe = evaluation.PipelinedEvaluator(net, B, T, depth=depth)
    best = None
    for rep in range(3):
        pipe.begin()
        for i in range(6):
            pipe.step(sets[i & 1][0], L, ld, sets[i & 1][1])
        pipe.finish(reduce=False)
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        pipe.begin()
        K = 60
        for i in range(K):
            pipe.step(sets[i & 1][0], L, ld, sets[i & 1][1])
        pipe.finish(reduce=False)
        e.record()
        torch.cuda.synchronize()
        ms = s.elapsed_time(e) / K
        best = ms if best is None else min(best, ms)
    print(f"depth {depth}: {best:.4f} ms per step")
