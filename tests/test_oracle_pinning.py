"""Pin the oracle restatements against outputs of the unmodified reference
(tests/golden/*.npz, written by oracle/make_golden.py in the build container)."""
import os

import numpy as np
import pytest
import torch

from artspeech_b200 import synthetic
from oracle import geometry as ogeo
from oracle import model as omodel
from oracle import reference_loader as ref


@pytest.fixture(scope="module")
def gm(golden_dir):
    return np.load(os.path.join(golden_dir, "model_small.npz"))


def test_seeded_weights_match_reference_init(gm):
    sd = omodel.make_state_dict(45, 10, seed=0)
    keys = sorted(sd.keys())
    assert keys == list(gm["keys"])
    sums = np.array([sd[k].double().sum().item() for k in keys])
    abssums = np.array([sd[k].double().abs().sum().item() for k in keys])
    assert np.array_equal(sums, gm["sums"])
    assert np.array_equal(abssums, gm["abssums"])
    assert sum(v.numel() for v in sd.values()) == 1739176


@pytest.mark.parametrize("affine", [False, True])
def test_model_forward_matches_reference(gm, affine):
    sd = omodel.make_state_dict(45, 10, seed=0)
    if affine:
        sd = omodel.randomize_affine(sd, seed=1)
    x = torch.from_numpy(gm["x"]); L = torch.from_numpy(gm["lengths"])
    sfx = "_affine" if affine else ""
    out32 = omodel.artspeech_forward(sd, x, L, torch.float32)
    out64 = omodel.artspeech_forward(sd, x, L, torch.float64)
    assert out32.shape == (3, 12, 10, 2, 50)
    assert np.abs(out32.numpy() - gm["out32" + sfx]).max() <= 2e-6
    assert np.abs(out64.numpy() - gm["out64" + sfx]).max() <= 1e-12


def test_gru_equations_match_stock(gm):
    sd = omodel.make_state_dict(45, 10, seed=0)
    x = torch.from_numpy(gm["x"]); L = torch.from_numpy(gm["lengths"])
    a = omodel.artspeech_forward(sd, x, L, torch.float64, rnn="equations")
    assert np.abs(a.numpy() - gm["out64"]).max() <= 1e-12


def test_lstm_equations_match_stock():
    sd = omodel.make_state_dict(45, 3, seed=2, cell="lstm")
    x = synthetic.tokens(2, 9, 45, seed=1); L = [9, 4]
    a = omodel.artspeech_forward(sd, x, L, torch.float64, rnn="equations")
    b = omodel.artspeech_forward(sd, x, L, torch.float64, rnn="stock")
    assert (a - b).abs().max() <= 1e-12


def test_padded_frames_are_constant(gm):
    """H3: padded GRU steps are zero, so padded frames equal decode(ReLU(b_linear))."""
    out = gm["out64"]
    assert np.abs(out[2, 5:] - out[2, 5]).max() == 0.0
    assert np.abs(out[1, 9:] - out[2, 11]).max() == 0.0


def test_unsorted_lengths_raise():
    sd = omodel.make_state_dict(45, 1, seed=0)
    x = synthetic.tokens(2, 6, 45)
    for rnn in ("stock", "equations"):
        with pytest.raises(RuntimeError):
            omodel.artspeech_forward(sd, x, [3, 6], rnn=rnn)


def test_p2cp_matches_reference(golden_dir):
    g = np.load(os.path.join(golden_dir, "p2cp.npz"))
    pred = torch.from_numpy(g["pred"]); tgt = torch.from_numpy(g["target"])
    got = ogeo.p2cp_distance(pred, tgt).numpy()
    assert np.max(np.abs(got - g["ref64"]) / g["ref64"]) <= 1e-12
    mean = ogeo.mean_p2cp(pred.transpose(-1, -2), tgt.transpose(-1, -2), "mean").item()
    assert abs(mean - g["mean64"]) <= 1e-14
    u2 = pred.transpose(-1, -2)[:4, :, :37]; v2 = tgt.transpose(-1, -2)[:4, :, 5:]
    rag = ogeo.mean_p2cp(u2, v2).numpy()
    assert np.max(np.abs(rag - g["ragged64"]) / g["ragged64"]) <= 1e-12
    med = ogeo.euclidean_distance(pred, tgt).numpy()
    assert np.max(np.abs(med - g["med64"])) <= 1e-14


def test_tv_matches_reference(golden_dir):
    g = np.load(os.path.join(golden_dir, "tv.npz"))
    frames = torch.from_numpy(g["frames"])
    val, idx, poc = ogeo.tract_variables_batch(frames, synthetic.TV_ART_INDEX)
    assert np.max(np.abs(val.numpy() - g["val64"])) <= 1e-15
    assert np.array_equal(poc.numpy(), g["poc64"])
    v2, i2, p2 = ogeo.tract_variables_batch_vectorized(frames, synthetic.TV_ART_INDEX)
    assert torch.equal(idx, i2) and torch.equal(poc, p2) and torch.equal(val, v2)
    # the tie frame (last): column-major first occurrence -> smallest j, then smallest i
    assert idx[-1, 0].tolist() == [0, 0]


def test_tv_tie_break_order():
    a1 = torch.tensor([[0., 0.], [1., 0.], [0., 0.]])
    a2 = torch.tensor([[5., 0.], [0., 1.], [1., 1.], [0., 1.]])
    v, i, j = ogeo.closest_pair(a1.double(), a2.double())
    assert (v, i, j) == (1.0, 0, 1)


def test_tv_dict_api(golden_dir):
    g = np.load(os.path.join(golden_dir, "tv.npz"))
    fr = torch.from_numpy(g["frames"])[3]
    d = {n: fr[k].T for k, n in enumerate(synthetic.TV_ARTICULATORS)}
    r = ogeo.calculate_vocal_tract_variables(d)
    assert r["LP"] is None and r["GLO"] is None
    assert abs(r["TBCD"]["value"] - g["val64"][3, 2]) <= 1e-15
    with pytest.raises(KeyError):
        bad = dict(d); bad.pop("tongue")
        ogeo.calculate_vocal_tract_variables(bad)


@pytest.mark.skipif(not ref.available(), reason="reference checkout absent")
def test_live_reference_agrees_with_oracle():
    m = ref.models()
    torch.manual_seed(4)
    net = m.ArtSpeech(30, 2).eval()
    sd = {k: v.detach().clone() for k, v in net.state_dict().items()}
    x = synthetic.tokens(4, 7, 30, seed=9); L = torch.tensor([7, 7, 3, 1])
    with torch.no_grad():
        want = net.double()(x, L)
    got = omodel.artspeech_forward(sd, x, L, torch.float64)
    assert (got - want).abs().max() <= 1e-12


def test_sentence_metrics_oracle_matches_reference_metrics_py(golden_dir):
    """oracle.geometry (euclidean_distance, pearsons_correlation incl. the x-target-mean quirk, p2cp)
    against per-sentence numbers produced by the reference's own metrics.py (oracle/make_golden.py)."""
    import numpy as np
    import torch
    from oracle import geometry as ogeo
    g = np.load(os.path.join(golden_dir, "sentence_metrics.npz"))
    pred, tgt = torch.from_numpy(g["pred"]), torch.from_numpy(g["target"])
    for b, L in enumerate(g["lengths"].tolist()):
        so, st = pred[b, :L].unsqueeze(0), tgt[b, :L].unsqueeze(0)
        p2cp = ogeo.p2cp_distance(so, st).mean(dim=1)[0].numpy()
        med = ogeo.euclidean_distance(so, st).mean(dim=1)[0].numpy()
        xc, yc = ogeo.pearsons_correlation(so, st)
        np.testing.assert_allclose(p2cp, g["per_sentence"][0, b], rtol=1e-12)
        np.testing.assert_allclose(med, g["per_sentence"][1, b], rtol=1e-12)
        np.testing.assert_allclose(xc.mean(dim=-1)[0].numpy(), g["per_sentence"][2, b], rtol=1e-10, atol=1e-12)
        np.testing.assert_allclose(yc.mean(dim=-1)[0].numpy(), g["per_sentence"][3, b], rtol=1e-10, atol=1e-12)


def test_simple_artspeech_constructor_and_oracle_match_reference(gm):
    """SimpleArtSpeech (models.py:53-96): the drop-in's constructor draws the reference's initial
    weights under the same seed, and the oracle restatement reproduces the reference's output."""
    from artspeech_b200 import models
    torch.manual_seed(0)
    net = models.SimpleArtSpeech(45, 10)
    sd = {k: v.detach() for k, v in net.state_dict().items()}
    sums = np.array([v.double().sum().item() for k, v in sorted(sd.items())])
    assert np.array_equal(sums, gm["simple_sums"])
    out = omodel.simple_artspeech_forward(sd, torch.from_numpy(gm["x"]))
    assert out.shape == gm["out_simple"].shape
    assert np.abs(out.numpy() - gm["out_simple"]).max() <= 2e-6


def test_critical_loss_oracle_matches_reference(golden_dir):
    """oracle.geometry.critical_loss against the reference's CriticalLoss (losses.py:23-99): value
    and gradient, with and without the injected upper incisor."""
    g = np.load(os.path.join(golden_dir, "critical_loss.npz"))
    arts = ["lower-lip", "pharynx", "soft-palate", "tongue", "upper-lip"]
    x = torch.from_numpy(g["x"]).double().requires_grad_(True)
    loss = ogeo.critical_loss(x, torch.from_numpy(g["ref"]), torch.from_numpy(g["mask"]),
                              ["LA", "TTCD", "TBCD", "VEL"], arts)
    loss.backward()
    assert abs(loss.item() - float(g["loss64"])) <= 1e-14
    assert np.abs(x.grad.numpy() - g["grad64"]).max() <= 1e-14
    assert abs(float(g["loss32"]) - float(g["loss64"])) / float(g["loss64"]) < 1e-4   # F2: fp32 cdist
    loss6 = ogeo.critical_loss(torch.from_numpy(g["x6"]), None, torch.from_numpy(g["mask"])[:, :2],
                               ["TTCD", "LA"], sorted(arts + ["upper-incisor"]))
    assert abs(loss6.item() - float(g["loss6"])) <= 1e-14
    if ref.available():
        CL = ref.critical_loss()
        live = CL(["TTCD", "LA"], sorted(arts + ["upper-incisor"]))(
            torch.from_numpy(g["x6"]).double(), torch.from_numpy(g["x6"]).double(), None,
            torch.from_numpy(g["mask"])[:, :2])
        assert abs(live.item() - float(g["loss6"])) <= 1e-14


def test_oracle_gradient_matches_reference(golden_dir):
    """Gradients through the oracle's from-the-equations forward (fp64) against the reference
    model's own autograd gradients (tests/golden/model_grad.npz)."""
    g = np.load(os.path.join(golden_dir, "model_grad.npz"))
    sd = {k: v.double().requires_grad_(True) for k, v in omodel.make_state_dict(45, 10, seed=0).items()}
    fwd = omodel.artspeech_forward.__wrapped__            # the undecorated (grad-enabled) function
    out = fwd(sd, torch.from_numpy(g["x"]), torch.from_numpy(g["lengths"]), torch.float64, rnn="equations")
    (out * torch.from_numpy(g["w"]).double()).sum().backward()
    keys = list(g["keys"])
    sums = np.array([sd[k].grad.sum().item() for k in keys])
    abssums = np.array([sd[k].grad.abs().sum().item() for k in keys])
    assert np.allclose(sums, g["sums64"], rtol=1e-9, atol=1e-10)
    assert np.allclose(abssums, g["abssums64"], rtol=1e-9, atol=1e-10)
    for k in ("linear.0.bias", "rnn.bias_hh_l0", "rnn.bias_ih_l1_reverse", "predictors.3.x_coords.bias"):
        assert np.allclose(sd[k].grad.numpy(), g["grad64:" + k], rtol=1e-9, atol=1e-11)


@pytest.mark.parametrize("cell", ["GRU", "LSTM"])
def test_pc_model_constructor_and_oracle_match_reference(golden_dir, cell):
    """PrincipalComponentsArtSpeech (principal_components/models/rnn.py:36-109): the drop-in's
    constructor draws the reference's initial weights under the same seed (same state_dict keys),
    and the oracle restatement reproduces the reference's fp32 / fp64 outputs."""
    from artspeech_b200 import principal_components as pc
    g = np.load(os.path.join(golden_dir, "pc_model.npz"))
    torch.manual_seed(0)
    net = pc.PrincipalComponentsArtSpeech(45, {"a": 4, "b": 5, "c": 3}, rnn=cell)
    sd = {k: v.detach() for k, v in net.state_dict().items()}
    assert sorted(sd.keys()) == list(g[f"keys_{cell}"]) and net.latent_size == 12
    sums = np.array([sd[k].double().sum().item() for k in sorted(sd.keys())])
    assert np.array_equal(sums, g[f"sums_{cell}"])
    x, L = torch.from_numpy(g["x"]), torch.from_numpy(g["lengths"])
    out32 = omodel.pc_artspeech_forward(sd, x, L, torch.float32)
    out64 = omodel.pc_artspeech_forward(sd, x, L, torch.float64, rnn="equations")
    assert np.abs(out32.numpy() - g[f"out32_{cell}"]).max() <= 2e-6
    assert np.abs(out64.numpy() - g[f"out64_{cell}"]).max() <= 1e-12
