"""Race hunt without compute-sanitizer (closed on this pool): every kernel of the path is deterministic by
construction, so repeated runs must agree BIT FOR BIT - alone, and with other streams' kernels running beside
them (three evaluation steps in flight).  Prints one line per case; exits 1 on the first difference."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from artspeech_b200 import evaluation, models, synthetic

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 30
bad = 0
for precision in ("fast", "fp32"):
    for cell in ("gru", "lstm"):
        net = models.ArtSpeech(45, 10, precision=precision, rnn=cell).cuda().eval()
        for (B, T) in [(1, 1), (3, 130), (41, 37), (256, 100), (700, 12)]:
            x = synthetic.tokens(B, T, 45, seed=B).cuda()
            L = [int(v) for v in synthetic.sorted_lengths(B, max(1, T // 3), T, seed=B)]
            with torch.no_grad():
                first = net(x, L).clone()
                diff = 0
                for _ in range(reps):
                    diff += int(not torch.equal(net(x, L), first))
            torch.cuda.synchronize()
            print(f"forward {precision:4s} {cell:4s} {B:4d} x {T:3d}: {diff} of {reps} runs differ", flush=True)
            bad += diff
    # three steps in flight: the accumulators of an evaluation, repeated
    net = models.ArtSpeech(45, 10, precision=precision).cuda().eval()
    B, T = 256, 100
    L = [int(v) for v in synthetic.sorted_lengths(B, 40, T, seed=1)]
    ld = torch.tensor(L, dtype=torch.int32, device="cuda")
    sets = [(synthetic.tokens(B, T, 45, seed=s).cuda(),
             synthetic.contours(B * T, 10, 50, seed=10 + s, device="cuda").view(B, T, 10, 2, 50)) for s in range(3)]
    pipe = evaluation.PipelinedEvaluator(net, B, T, depth=3)
    first, diff = None, 0
    for rep in range(reps):
        pipe.begin()
        for i in range(9):
            pipe.step(sets[i % 3][0], L, ld, sets[i % 3][1])
        total = pipe.finish(reduce=False).clone()
        torch.cuda.synchronize()
        if first is None:
            first = total
        diff += int(not torch.equal(total, first))
    print(f"pipelined evaluation {precision}: {diff} of {reps} evaluations differ; frames {float(first[-1]):.0f}", flush=True)
    bad += diff
print("OK" if bad == 0 else f"{bad} DIFFERENCES")
sys.exit(1 if bad else 0)
