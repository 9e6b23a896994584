"""Generate tests/golden/*.npz by running the UNMODIFIED reference (build container only).

TEST INFRASTRUCTURE ONLY.  Run as ``python -m oracle.make_golden`` from the repo root
where /root/reference exists.  The fixtures hold seeded synthetic inputs and the
reference's outputs on them, so that the GPU box (which has no reference checkout)
can pin both the oracle restatements and the CUDA path.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from artspeech_b200 import synthetic  # noqa: E402
from oracle import model as omodel  # noqa: E402
from oracle import reference_loader as ref  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")


def golden_model():
    m = ref.models()
    torch.manual_seed(0)
    net = m.ArtSpeech(45, 10).eval()
    sd = {k: v.detach().clone() for k, v in net.state_dict().items()}
    # per-key fp64 sums pin oracle.model.make_state_dict(seed=0) to the reference init
    keys = sorted(sd.keys())
    sums = np.array([sd[k].double().sum().item() for k in keys])
    abssums = np.array([sd[k].double().abs().sum().item() for k in keys])

    x = synthetic.tokens(3, 12, 45, seed=3)
    lengths = torch.tensor([12, 9, 5], dtype=torch.int32)
    with torch.no_grad():
        out32 = net(x, lengths)
        out64 = net.double()(x, lengths)
    net.float()

    # second case: LayerNorm affine terms perturbed so they are exercised
    sd2 = omodel.randomize_affine(sd, seed=1)
    net2 = m.ArtSpeech(45, 10).eval()
    net2.load_state_dict(sd2)
    with torch.no_grad():
        out32_b = net2(x, lengths)
        out64_b = net2.double()(x, lengths)

    # SimpleArtSpeech (models.py:53-96) shares the predictors
    torch.manual_seed(0)
    simple = m.SimpleArtSpeech(45, 10).eval()
    with torch.no_grad():
        out_simple = simple(x, lengths)
    simple_sums = np.array([v.double().sum().item() for k, v in sorted(simple.state_dict().items())])

    np.savez_compressed(
        os.path.join(GOLDEN, "model_small.npz"),
        keys=np.array(keys), sums=sums, abssums=abssums,
        x=x.numpy(), lengths=lengths.numpy(),
        out32=out32.numpy(), out64=out64.numpy(),
        out32_affine=out32_b.numpy(), out64_affine=out64_b.numpy(),
        out_simple=out_simple.numpy(), simple_sums=simple_sums,
    )
    print("model_small:", tuple(out32.shape), "params", sum(v.numel() for v in sd.values()))


def golden_predictor():
    """Standalone ArticulatorPredictor (models.py:7-33; instantiated on its own by
    transformer/models.py:337-339) of the unmodified reference: seed-0 instances with LayerNorm affines
    perturbed, fp32 and fp64 outputs, per-key parameter sums."""
    m = ref.models()
    out = {}
    for tag, (c, d, shape) in {"a": (64, 50, (5, 13)), "b": (256, 30, (70,)), "c": (128, 50, (2, 3, 4))}.items():
        torch.manual_seed(0)
        net = m.ArticulatorPredictor(c, d).eval()
        g = torch.Generator().manual_seed(17)
        with torch.no_grad():
            for k, v in net.state_dict().items():      # exercise the affine terms
                if k.startswith("linear.") and v.dim() == 1 and k.split(".")[1] in ("0", "3", "6"):
                    v.add_(0.3 * torch.randn(v.shape, generator=g))
        sd = net.state_dict()
        keys = sorted(sd.keys())
        x = torch.randn(*shape, c, generator=torch.Generator().manual_seed(18))
        with torch.no_grad():
            o32 = net(x)
            o64 = net.double()(x.double())
        out.update({f"x_{tag}": x.numpy(), f"out32_{tag}": o32.numpy(), f"out64_{tag}": o64.numpy(),
                    f"keys_{tag}": np.array(keys),
                    f"sums_{tag}": np.array([sd[k].double().sum().item() for k in keys])})
    np.savez_compressed(os.path.join(GOLDEN, "predictor.npz"), **out)
    print("predictor:", out["out32_a"].shape, out["out32_b"].shape, out["out32_c"].shape)


def golden_model_grad():
    """Gradients of the unmodified reference ArtSpeech (fp32 and fp64) for a seeded linear functional
    of its output: per-parameter sums / abs-sums plus the full gradient of the small tensors."""
    m = ref.models()
    torch.manual_seed(0)
    net = m.ArtSpeech(45, 10).train()          # dropout = 0: train mode == eval mode numerically
    x = synthetic.tokens(3, 12, 45, seed=3)
    lengths = torch.tensor([12, 9, 5], dtype=torch.int32)
    w = torch.randn(3, 12, 10, 2, 50, generator=torch.Generator().manual_seed(11))
    out = {}
    for name, dt in (("32", torch.float32), ("64", torch.float64)):
        net = net.to(dt)
        net.zero_grad()
        (net(x, lengths) * w.to(dt)).sum().backward()
        keys = sorted(k for k, _ in net.named_parameters())
        g = {k: p.grad.detach().double() for k, p in net.named_parameters()}
        out["sums" + name] = np.array([g[k].sum().item() for k in keys])
        out["abssums" + name] = np.array([g[k].abs().sum().item() for k in keys])
        for k in ("linear.0.bias", "rnn.bias_hh_l0", "rnn.bias_ih_l1_reverse", "predictors.3.x_coords.bias"):
            out[f"grad{name}:{k}"] = g[k].numpy()
    np.savez_compressed(os.path.join(GOLDEN, "model_grad.npz"), keys=np.array(keys), x=x.numpy(),
                        lengths=lengths.numpy(), w=w.numpy(), **out)
    print("model_grad:", len(keys), "parameters, |grad| total", out["abssums64"].sum())


def golden_pc_model():
    """PrincipalComponentsArtSpeech of the unmodified reference (class statements cut out with ast):
    GRU and LSTM instances under seed 0, fp32 and fp64 outputs, per-key parameter sums."""
    PC, RNNType = ref.principal_components_models()
    x = synthetic.tokens(3, 12, 45, seed=3)
    lengths = torch.tensor([12, 9, 5], dtype=torch.int32)
    out = {}
    for cell in ("GRU", "LSTM"):
        torch.manual_seed(0)
        net = PC(45, {"a": 4, "b": 5, "c": 3}, rnn=RNNType[cell]).eval()
        sd = net.state_dict()
        out[f"keys_{cell}"] = np.array(sorted(sd.keys()))
        out[f"sums_{cell}"] = np.array([sd[k].double().sum().item() for k in sorted(sd.keys())])
        with torch.no_grad():
            out[f"out32_{cell}"] = net(x, lengths).numpy()
            out[f"out64_{cell}"] = net.double()(x, lengths).numpy()
    np.savez_compressed(os.path.join(GOLDEN, "pc_model.npz"), x=x.numpy(), lengths=lengths.numpy(), **out)
    print("pc_model:", out["out32_GRU"].shape, float(np.abs(out["out32_GRU"]).max()))


def golden_p2cp():
    pm = ref.p2a_metrics()
    tgt = synthetic.contours(24, 10, 50, seed=5)
    pred = synthetic.perturbed(tgt, seed=6)
    fn = pm.MeanP2CPDistance(reduction="none")
    u, v = pred.transpose(-1, -2), tgt.transpose(-1, -2)
    with torch.no_grad():
        ref64 = fn(u.double(), v.double())
        ref32 = fn(u, v)
        mean64 = pm.MeanP2CPDistance(reduction="mean")(u.double(), v.double())
        # ragged point counts (n != m) through the generic (*, N, 2) entry
        u2 = u[:4, :, :37, :].contiguous()
        v2 = v[:4, :, 5:, :].contiguous()
        ragged64 = fn(u2.double(), v2.double())
        med64 = pm.EuclideanDistance(reduction="none")(pred.double(), tgt.double()).mean(dim=-1)
    np.savez_compressed(
        os.path.join(GOLDEN, "p2cp.npz"),
        pred=pred.numpy(), target=tgt.numpy(), ref64=ref64.numpy(), ref32=ref32.numpy(),
        mean64=mean64.numpy(), ragged64=ragged64.numpy(), med64=med64.numpy(),
    )
    rel = ((ref32.double() - ref64).abs() / ref64).max().item()
    print("p2cp:", tuple(ref64.shape), "reference fp32-vs-fp64 max rel", rel)


def golden_sentence_metrics():
    """run_test's per-sentence numbers (evaluation.py:68-84) from the reference's own metrics.py:
    p2cp / euclidean means over the sentence and pearsons_correlation with the sentence as batch."""
    rm = ref.root_metrics()
    B, T, N = 3, 17, 10
    lengths = [17, 11, 1]
    tgt = synthetic.contours(B * T, N, 50, seed=15).view(B, T, N, 2, 50)
    pred = synthetic.perturbed(tgt.view(B * T, N, 2, 50), seed=16).view(B, T, N, 2, 50)
    # a smooth drift over time so that the correlations are not all ~1
    drift = 0.02 * torch.sin(torch.arange(T, dtype=torch.float32) * 0.7).view(1, T, 1, 1, 1)
    pred = pred + drift * torch.linspace(-1, 1, 50).view(1, 1, 1, 1, 50)
    out = np.zeros((4, B, N))
    with torch.no_grad():
        for b, L in enumerate(lengths):
            so = pred[b, :L].unsqueeze(0).double()
            st = tgt[b, :L].unsqueeze(0).double()
            out[0, b] = rm.p2cp_distance(so, st).mean(dim=1)[0].numpy()
            out[1, b] = rm.euclidean_distance(so, st).mean(dim=1)[0].numpy()
            xc, yc = rm.pearsons_correlation(so, st)
            out[2, b] = xc.mean(dim=-1)[0].numpy()
            out[3, b] = yc.mean(dim=-1)[0].numpy()
    np.savez_compressed(os.path.join(GOLDEN, "sentence_metrics.npz"), pred=pred.numpy(), target=tgt.numpy(),
                        lengths=np.array(lengths), per_sentence=out)
    print("sentence_metrics:", out.shape, "x_corr range", out[2].min(), out[2].max())


def golden_tv():
    tv = ref.tract_variables()
    frames = synthetic.tv_frames(24, seed=7)
    # a degenerate frame with exact ties: every contour collapsed onto lattice points
    tie = frames[0].clone()
    tie[:, 0, :] = (torch.arange(50) % 2).float() * 0.25 + 0.25
    tie[:, 1, :] = 0.5
    tie[synthetic.TV_ART_INDEX["upper-lip"], 1, :] = 0.75
    tie[synthetic.TV_ART_INDEX["upper-incisor"], 1, :] = 0.75
    tie[synthetic.TV_ART_INDEX["pharynx"], 1, :] = 0.75
    frames = torch.cat([frames, tie[None]], dim=0)
    names = synthetic.TV_ARTICULATORS
    F = frames.shape[0]
    val = np.zeros((F, 4)); poc = np.zeros((F, 4, 2, 2))
    val32 = np.zeros((F, 4))
    for f in range(F):
        d64 = {n: frames[f, k].double().T for k, n in enumerate(names)}
        d32 = {n: frames[f, k].T for k, n in enumerate(names)}
        r64 = tv.calculate_vocal_tract_variables(d64)
        r32 = tv.calculate_vocal_tract_variables(d32)
        for k, name in enumerate(("LA", "TTCD", "TBCD", "VEL")):
            val[f, k] = r64[name]["value"]
            val32[f, k] = r32[name]["value"]
            poc[f, k, 0] = r64[name]["poc_1"].numpy()
            poc[f, k, 1] = r64[name]["poc_2"].numpy()
        assert all(r64[n] is None for n in ("LP", "TTCL", "TBCL", "GLO"))
    np.savez_compressed(os.path.join(GOLDEN, "tv.npz"), frames=frames.numpy(),
                        val64=val, poc64=poc, val32=val32)
    print("tv:", F, "frames; reference fp32-vs-fp64 max rel",
          np.max(np.abs(val32 - val) / np.maximum(val, 1e-30)))


def golden_critical_loss():
    """CriticalLoss of the unmodified reference (class statement only, see reference_loader) on
    seeded contours: loss value and its gradient, in fp64 and fp32."""
    CL = ref.critical_loss()
    arts = ["lower-lip", "pharynx", "soft-palate", "tongue", "upper-lip"]   # upper incisor injected
    B, T = 3, 7
    x = synthetic.contours(B * T, 5, 50, seed=21).view(B, T, 5, 2, 50)
    refarr = synthetic.UPPER_INCISOR_REF.view(1, 1, 1, 2, 50).expand(B, T, 1, 2, 50).contiguous()
    g = torch.Generator().manual_seed(5)
    mask = (torch.rand(B, 4, T, generator=g) < 0.4).to(torch.int64)
    mask[0, 0, 0] = 1
    out = {}
    for name, dt in (("64", torch.float64), ("32", torch.float32)):
        xi = x.to(dt).clone().requires_grad_(True)
        loss = CL(["LA", "TTCD", "TBCD", "VEL"], list(arts))(xi, xi.detach(), refarr.to(dt), mask)
        loss.backward()
        out["loss" + name] = np.array(loss.item())
        out["grad" + name] = xi.grad.numpy()
    # a two-TV instance whose articulator list already holds the upper incisor (no injection)
    arts6 = sorted(arts + ["upper-incisor"])
    x6 = synthetic.contours(B * T, 6, 50, seed=22).view(B, T, 6, 2, 50).double()
    loss6 = CL(["TTCD", "LA"], list(arts6))(x6, x6, None, mask[:, :2])
    np.savez_compressed(os.path.join(GOLDEN, "critical_loss.npz"), x=x.numpy(), ref=refarr.numpy(), mask=mask.numpy(),
                        x6=x6.float().numpy(), loss6=np.array(loss6.item()), **out)
    print("critical_loss:", out["loss64"], out["loss32"], float(loss6))


def golden_tv_driver():
    """tract_variables.csv written by the reference's own frame driver (__init__.py:201-297) for two
    short sentences (float64 copies of fp32 contours, so its cdist is the exact one)."""
    import shutil
    import tempfile
    drv = ref.tract_variables_driver()
    names = synthetic.TV_ARTICULATORS
    frames = synthetic.tv_frames(2 * 6, seed=13).view(2, 6, len(names), 2, 50)
    pred = frames + 0.01 * torch.randn(frames.shape, generator=torch.Generator().manual_seed(3))
    pred = pred.float()
    lengths = [6, 4]
    frame_ids = [["%04d" % (10 + k) for k in range(6)], ["%04d" % (200 + k) for k in range(4)]]
    phonemes = [list("abcdef"), list("wxyz")]
    tmp = tempfile.mkdtemp()
    drv(["S1", "S2"], frame_ids, pred.double(), frames.double(), lengths, phonemes, list(names), tmp)
    for sid in ("S1", "S2"):
        shutil.copy(os.path.join(tmp, sid, "tract_variables.csv"), os.path.join(GOLDEN, f"tv_driver_{sid}.csv"))
    shutil.rmtree(tmp)
    np.savez_compressed(os.path.join(GOLDEN, "tv_driver.npz"), pred=pred.numpy(), target=frames.numpy())
    print("tv_driver: 2 sentences,", sum(lengths), "frames")


def golden_run_test():
    """``info`` returned by the reference's own run_test (evaluation.py:17-160) on the reference
    ArtSpeech (CPU, seed-0 weights) over a synthetic two-batch loader, plus one of the
    tract_variables.csv / contour files it wrote."""
    import json
    import shutil
    import tempfile
    rt = ref.run_test()
    m = ref.models()
    pm = ref.p2a_metrics()
    torch.manual_seed(0)
    net = m.ArtSpeech(45, 10).eval()
    tmp = tempfile.mkdtemp()
    info = rt(3, net, synthetic.loader(), pm.EuclideanDistance("none"), tmp, list(synthetic.ARTICULATORS),
              device=torch.device("cpu"))
    info = {k: (float(v) if not isinstance(v, dict) else {a: float(b) for a, b in v.items()}) for k, v in info.items()}
    with open(os.path.join(GOLDEN, "run_test_info.json"), "w") as f:
        json.dump(info, f, indent=1, sort_keys=True)
    shutil.copy(os.path.join(tmp, "3", "S1_2", "tract_variables.csv"), os.path.join(GOLDEN, "run_test_S1_2_tv.csv"))
    files = sorted(os.listdir(os.path.join(tmp, "3", "S0_0", "contours")))
    with open(os.path.join(GOLDEN, "run_test_S0_0_files.txt"), "w") as f:
        f.write("\n".join(files) + "\n")
    shutil.rmtree(tmp)
    print("run_test: loss", info["loss"], "tongue", info["tongue"])


def golden_area():
    """The parts of area_function.py that run without GEOS: grid, mid_point, argmatrix, rotate and
    area_function (a9, a11).  a10 / a12 need shapely and stay unpinned (oracle/area.py header)."""
    import math
    from oracle import area as oarea
    af = ref.area_function()
    grids = {}
    cases = {"default": synthetic.AREA_GRID_ARGS,
             "rotated": dict(center=(0.45, 0.55), theta_rad=0.2, omega_rad=-0.15, linear_step=0.02,
                             polar_step_rad=math.pi / 30)}
    for name, kw in cases.items():
        grids[name] = af.build_semipolar_grid(np.array(kw["center"]), kw["theta_rad"], kw["omega_rad"],
                                              kw["linear_step"], kw["polar_step_rad"])
    # wall pairs for area_function(): intersections produced by the restatement on synthetic walls
    wi, we = synthetic.walls(3, seed=11)
    out = {}
    for f in range(3):
        ii, ee = oarea.intersect_semipolar_grid(wi[f].numpy(), we[f].numpy(), grids["default"])
        dists, fx = af.area_function(ii, ee)
        dists_b, fx_b = af.area_function(ii, ee, alpha=2.5, beta=1.5)
        out.update({f"int_{f}": ii, f"ext_{f}": ee, f"dists_{f}": dists, f"fx_{f}": fx,
                    f"dists_b_{f}": dists_b, f"fx_b_{f}": fx_b})
    rng = np.random.default_rng(3)
    p = rng.random((16, 2, 2))
    mids = np.array([af.mid_point(a, b) for a, b in p])
    mtx = rng.random((5, 7)); mtx[3, 2] = mtx[1, 4] = -1.0   # tie: first in row-major order wins
    arg = np.array([af.argmatrix(mtx, "min"), af.argmatrix(mtx, "max")])
    rot = np.array([af.rotate(np.array([0.3, -0.4]), a) for a in (0.0, 0.7, -2.1)])
    np.savez_compressed(os.path.join(GOLDEN, "area.npz"), grid_default=grids["default"],
                        grid_rotated=grids["rotated"], wall_int=wi.numpy(), wall_ext=we.numpy(),
                        mid_in=p, mid_out=mids, arg_in=mtx, arg_out=arg, rot_out=rot, **out)
    print("area: grids", grids["default"].shape, grids["rotated"].shape, "K", [len(out[f"fx_{f}"]) for f in range(3)])


def main(out_dir=None):
    """Run every generator in ONE process, writing into ``out_dir`` (default tests/golden)."""
    global GOLDEN
    assert ref.available(), "reference checkout not found"
    keep = GOLDEN
    if out_dir is not None:
        GOLDEN = out_dir
    os.makedirs(GOLDEN, exist_ok=True)
    try:
        _run_all()
    finally:
        GOLDEN = keep


def _run_all():
    golden_model()
    golden_predictor()
    golden_model_grad()
    golden_pc_model()
    golden_p2cp()
    golden_tv()
    golden_sentence_metrics()
    golden_area()
    golden_run_test()
    golden_tv_driver()
    golden_critical_loss()


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else None)
