"""us per recurrent step (batch 256 x 100 and 64 x 1024) and parity for every tools/_var_rnn_*.so + the library."""
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
    from oracle import model as omodel
    sd = omodel.randomize_affine(omodel.make_state_dict(45, 10, seed=0), seed=3)
    net = models.ArtSpeech(45, 10, precision="fast")
    net.load_state_dict(sd)
    net = net.cuda().eval()
    x = synthetic.tokens(8, 60, 45, seed=5)
    L = synthetic.sorted_lengths(8, 20, 60, seed=5)
    with torch.no_grad():
        out = net(x.cuda(), L)
    want = omodel.artspeech_forward(sd, x, L, torch.float32)
    res = {"lib": os.path.basename(sys.argv[2]), "err": float((out[:, :want.shape[1]].cpu() - want).abs().max())}
    for (B, T) in [(256, 100), (64, 1024)]:
        xx = synthetic.tokens(B, T, 45, seed=1).cuda()
        Lf = [T] * B
        ld = torch.tensor(Lf, dtype=torch.int32, device="cuda")
        with torch.no_grad():
            for _ in range(3):
                net(xx, Lf, lengths_dev=ld)
            net.profile(True)
            net.read_profile(reset=True)
            for _ in range(10):
                net(xx, Lf, lengths_dev=ld)
            torch.cuda.synchronize()
            prof = net.read_profile(reset=True)
            net.profile(False)
        res[f"us_per_step_{B}x{T}"] = round(1e3 * prof["gru_l0"][0] / prof["gru_l0"][1] / T, 4)
    print(json.dumps(res))
else:
    for lib in sorted(glob.glob("tools/_var_rnn_*.so")) + ["artspeech_b200/libartspeech_b200.so"]:
        if "trace" in lib:
            continue
        subprocess.run([sys.executable, __file__, "--one", lib], timeout=200)
