"""Per-stage forward times at configs[1] for every tools/_var_*.so (one process each)."""
import glob
import json
import os
import subprocess
import sys
if len(sys.argv) > 1 and sys.argv[1] == "--one":
    sys.path.insert(0, ".")
    import torch
    from artspeech_b200 import _lib
    _lib.LIB_PATH = sys.argv[2]
    from artspeech_b200 import models, synthetic
    B, T = 256, 100
    net = models.ArtSpeech(45, 10, precision="fast").cuda().eval()
    x = synthetic.tokens(B, T, 45, seed=5).cuda()
    L = [T] * B
    ld = torch.tensor(L, dtype=torch.int32, device="cuda")
    with torch.no_grad():
        for _ in range(5):
            net(x, L, lengths_dev=ld)
        net.profile(True)
        for _ in range(30):
            net(x, L, lengths_dev=ld)
        torch.cuda.synchronize()
    prof = {k: round(v[0] / max(1, v[1]), 4) for k, v in net.read_profile().items()}
    print(json.dumps({"lib": os.path.basename(sys.argv[2]), **prof}))
else:
    for lib in sorted(glob.glob("tools/_var_*.so")) + ["artspeech_b200/libartspeech_b200.so"]:
        subprocess.run([sys.executable, __file__, "--one", lib], timeout=120)
