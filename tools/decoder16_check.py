"""fp16-operand tcgen05 decoder (decoder_umma16.cu) against the tf32 one and the fp32 oracle: error and time."""
import json
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from artspeech_b200 import models, synthetic
from oracle import model as omodel

res = {}
for gamma in (1.0, 2.0):
    sd = omodel.randomize_affine(omodel.make_state_dict(45, 10, seed=0), seed=1)
    if gamma != 1.0:
        for k in list(sd):
            part = k.split(".")
            if k.startswith("predictors.") and part[2] == "linear" and part[3] in ("0", "3", "6") and k.endswith("weight"):
                sd[k] = sd[k] * gamma
    net = models.ArtSpeech(45, 10, precision="fast")
    net.load_state_dict(sd)
    net = net.cuda().eval()
    B, T = 256, 100
    x = synthetic.tokens(B, T, 45, seed=0)
    L = [T] * B
    want = omodel.artspeech_forward(sd, x[:16], L[:16], torch.float32)
    for impl in ("tcgen05", "tcgen05-tf32", "mma"):
        net.set_decoder(impl)
        with torch.no_grad():
            out = net(x.cuda(), L)
            torch.cuda.synchronize()
            err = float((out[:16].cpu() - want).abs().max())
            nan = bool(torch.isnan(out).any())
            net.profile(True); net.read_profile(reset=True)
            for _ in range(10):
                torch.cuda._sleep(400000)
                net(x.cuda(), L)
            torch.cuda.synchronize()
            prof = net.read_profile(reset=True)
            net.profile(False)
        res[f"gamma{gamma}:{impl}"] = {"max_abs_err": err, "nan": nan, "decoder_ms": prof["decoder"][0] / prof["decoder"][1]}
print(json.dumps(res, indent=1))
