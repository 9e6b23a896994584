"""GPU parity of the drop-in PrincipalComponentsArtSpeech (principal_components/models/rnn.py:36-109)
against the reference's golden outputs (tests/golden/pc_model.npz) and the oracle."""
import os

import numpy as np
import pytest
import torch

from artspeech_b200 import synthetic

pytestmark = pytest.mark.gpu

# The components are tanh outputs in (-1, 1): twice the range of the contours' sigmoid and four times its
# slope, so the fast-mode gate of the contours (2e-3 in a unit range) corresponds to 4e-3 here (measured
# 1.7e-3 ... 3.0e-3); fp32 mode keeps the 1e-5 gate (measured 5e-6 ... 6e-6).
TOL = {"fast": 4e-3, "fp32": 1e-5}


@pytest.fixture(scope="module")
def mods():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from artspeech_b200 import principal_components as pc
    from oracle import model as omodel
    return pc, omodel


@pytest.mark.parametrize("cell,precision", [("GRU", "fast"), ("GRU", "fp32"), ("LSTM", "fast"), ("LSTM", "fp32")])
def test_pc_model_golden(mods, golden_dir, cell, precision):
    pc, omodel = mods
    g = np.load(os.path.join(golden_dir, "pc_model.npz"))
    torch.manual_seed(0)
    net = pc.PrincipalComponentsArtSpeech(45, {"a": 4, "b": 5, "c": 3}, rnn=cell, precision=precision).cuda().eval()
    out = net(torch.from_numpy(g["x"]).cuda(), torch.from_numpy(g["lengths"]))
    want = g[f"out32_{cell}"]
    assert out.shape == want.shape and out.dtype == torch.float32 and out.is_contiguous()
    err = np.abs(out.cpu().numpy() - want).max()
    print(f"pc_model[{cell}, {precision}] max-abs vs reference {err:.3e}")
    assert err <= TOL[precision]


@pytest.mark.parametrize("precision", ["fast", "fp32"])
def test_pc_model_batch_vs_oracle(mods, precision):
    """Partly filled tiles, variable lengths, perturbed LayerNorm affines, a latent size that is not a
    multiple of the MMA tile; padded frames equal the constant head(ReLU(b_linear)) row."""
    pc, omodel = mods
    torch.manual_seed(3)
    net = pc.PrincipalComponentsArtSpeech(45, {"a": 7, "b": 6}, precision=precision)
    with torch.no_grad():
        for k, p in net.named_parameters():
            if "predictor.linear" in k and p.dim() == 1:
                p.add_(0.3 * torch.randn_like(p))
    sd = {k: v.detach().clone() for k, v in net.state_dict().items()}
    net = net.cuda().eval()
    B, T = 37, 61
    x = synthetic.tokens(B, T, 45, seed=5)
    L = synthetic.sorted_lengths(B, 7, 53, seed=5)
    out = net(x.cuda(), L)
    want = omodel.pc_artspeech_forward(sd, x, L, torch.float32)
    assert out.shape == want.shape == (B, 53, 13)
    err = (out.cpu() - want).abs().max().item()
    print(f"pc_model_batch[{precision}] max-abs {err:.3e}")
    assert err <= TOL[precision]
    pad = out[B - 1, int(L[B - 1]):]
    assert (pad - pad[0]).abs().max().item() == 0.0
    with pytest.raises(RuntimeError):
        net(x.cuda(), list(reversed(L.tolist())))          # unsorted lengths, like pack_padded_sequence
