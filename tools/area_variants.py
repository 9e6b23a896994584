"""Time asb_area_function from several builds of the library (tools/_var_*.so), one process each."""
import glob, json, os, subprocess, sys
if len(sys.argv) > 1 and sys.argv[1] == "--one":
    sys.path.insert(0, ".")
    import torch
    from artspeech_b200 import _lib, synthetic
    _lib.LIB_PATH = sys.argv[2]
    import artspeech_b200.area_function as AF
    F = 1 << 19
    wi, we = synthetic.walls(4096, seed=2)
    wi = wi.cuda().repeat(F // 4096, 1, 1).contiguous(); we = we.cuda().repeat(F // 4096, 1, 1).contiguous()
    grid = AF.DeviceGrid(AF.build_semipolar_grid(**synthetic.AREA_GRID_ARGS), wi.device)
    for _ in range(3):
        AF.area_function_batch(wi, we, grid)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        AF.area_function_batch(wi, we, grid)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print(json.dumps({"lib": os.path.basename(sys.argv[2]), "ms": ms, "frames_per_s": F / ms * 1e3}))
else:
    for lib in sorted(glob.glob("tools/_var_*.so")):
        subprocess.run([sys.executable, __file__, "--one", lib])
