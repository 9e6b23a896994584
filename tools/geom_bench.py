"""Microbenchmark of the geometry kernels (CUDA events, inputs larger than L2)."""
import json
import sys

import torch

sys.path.insert(0, ".")
from artspeech_b200 import _lib, synthetic  # noqa: E402
import artspeech_b200.metrics as M  # noqa: E402
import artspeech_b200.tract_variables as TV  # noqa: E402
import artspeech_b200.area_function as AF  # noqa: E402


def timeit(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(iters + 1)]
    ev[0].record()
    for i in range(iters):
        fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    ts = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(iters))
    return ts[len(ts) // 2], ts[0]


def main():
    F = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
    base = synthetic.tv_frames(4096, seed=1).cuda()
    reps = F // 4096
    tgt = base.repeat(reps, 1, 1, 1).contiguous()
    pred = (tgt + 0.01 * torch.randn_like(tgt)).contiguous()
    p10 = pred[:, :10].contiguous(); t10 = tgt[:, :10].contiguous()
    res = {"frames": F}
    med, best = timeit(lambda: M.p2cp_distance(p10, t10))
    res["p2cp_ms"] = med; res["p2cp_frames_per_s"] = F / med * 1e3
    res["p2cp_GBs"] = F * 8040 / med * 1e-6
    med, best = timeit(lambda: (TV.tract_variables_batch(pred, synthetic.TV_ARTICULATORS),
                                TV.tract_variables_batch(tgt, synthetic.TV_ARTICULATORS)))
    res["tv_ms"] = med; res["tv_frames_per_s"] = F / med * 1e3
    res["tv_GBs"] = F * 5024 / med * 1e-6
    wi, we = synthetic.walls(4096, seed=2)
    wi = wi.cuda().repeat(reps, 1, 1).contiguous(); we = we.cuda().repeat(reps, 1, 1).contiguous()
    grid = AF.DeviceGrid(AF.build_semipolar_grid(**synthetic.AREA_GRID_ARGS), wi.device)
    med, best = timeit(lambda: AF.area_function_batch(wi, we, grid))
    res["area_ms"] = med; res["area_frames_per_s"] = F / med * 1e3
    res["area_GBs"] = F * 2404 / med * 1e-6
    res["launches"] = _lib.launch_count()
    print(json.dumps(res))


if __name__ == "__main__":
    main()
