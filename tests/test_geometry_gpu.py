"""GPU parity of the geometry kernels (through the C ABI) against the fp64 oracle and the
reference's golden outputs.  Gates (BASELINE.json north_star): values within 1e-5 relative,
indices bit-exact."""
import os

import numpy as np
import pytest
import torch

from artspeech_b200 import synthetic

pytestmark = pytest.mark.gpu

REL = 1e-5


@pytest.fixture(scope="module")
def asb():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import artspeech_b200.metrics as M
    import artspeech_b200.tract_variables as TV
    from artspeech_b200 import _lib
    _lib.lib()
    return M, TV


def _rel(a, b):
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-30)))


def test_p2cp_golden(asb, golden_dir):
    M, _ = asb
    g = np.load(os.path.join(golden_dir, "p2cp.npz"))
    pred = torch.from_numpy(g["pred"]).cuda(); tgt = torch.from_numpy(g["target"]).cuda()
    got = M.p2cp_distance(pred, tgt)
    assert got.shape == (24, 10) and got.is_cuda
    assert _rel(got.cpu(), g["ref64"]) <= REL
    # CPU tensors in -> CPU tensor out, as the reference's evaluation loop calls it
    got_cpu = M.p2cp_distance(pred.cpu()[None], tgt.cpu()[None])
    assert not got_cpu.is_cuda and got_cpu.shape == (1, 24, 10)
    assert _rel(got_cpu[0], g["ref64"]) <= REL
    mean = M.MeanP2CPDistance("mean")(pred.transpose(-1, -2), tgt.transpose(-1, -2))
    assert abs(mean.item() - float(g["mean64"])) <= REL * float(g["mean64"])
    # ragged n != m through the generic (*, N, 2) entry
    u2 = pred.transpose(-1, -2)[:4, :, :37].contiguous(); v2 = tgt.transpose(-1, -2)[:4, :, 5:].contiguous()
    rag = M.MeanP2CPDistance("none")(u2, v2)
    assert _rel(rag.cpu(), g["ragged64"]) <= REL
    med = M.euclidean_distance(pred, tgt)
    assert _rel(med.cpu(), g["med64"]) <= REL


@pytest.mark.parametrize("items", [1, 2, 127, 128, 129, 1000, 148 * 128 * 2 + 77])
def test_p2cp_vs_oracle_sizes(asb, items):
    from oracle import geometry as og
    M, _ = asb
    tgt = synthetic.contours(items, 1, 50, seed=items)[:, 0]
    pred = synthetic.perturbed(tgt, seed=items + 1)
    got = M.p2cp_distance(pred.cuda(), tgt.cuda()).cpu()
    want = og.p2cp_distance(pred, tgt)
    assert got.shape == want.shape
    assert _rel(got, want) <= REL
    # fast path and generic path agree (interleaved copies take the generic kernel)
    gen = M.MeanP2CPDistance("none")(pred.transpose(-1, -2).contiguous().cuda(),
                                     tgt.transpose(-1, -2).contiguous().cuda()).cpu()
    assert _rel(gen, want) <= REL


def test_p2cp_properties(asb):
    M, _ = asb
    c = synthetic.contours(4096, 1, 50, seed=11)[:, 0].cuda()
    assert float(M.p2cp_distance(c, c).abs().max()) == 0.0           # identical contours -> 0
    d = synthetic.perturbed(c.cpu(), seed=12).cuda()
    a = M.p2cp_distance(c, d); b = M.p2cp_distance(d, c)
    assert torch.allclose(a, b, rtol=1e-6, atol=0)                   # symmetric in its arguments
    shift = torch.tensor([0.125, -0.25], device="cuda").view(1, 2, 1)
    e = M.p2cp_distance(c + shift, d + shift)                        # translation invariant
    assert torch.allclose(a, e, rtol=2e-5, atol=0)
    assert M.p2cp_distance(c[:0], d[:0]).shape == (0,)               # empty input


def test_tv_golden(asb, golden_dir):
    _, TV = asb
    from oracle import geometry as og
    g = np.load(os.path.join(golden_dir, "tv.npz"))
    frames = torch.from_numpy(g["frames"])
    val, idx, poc = TV.tract_variables_batch(frames.cuda(), synthetic.TV_ARTICULATORS)
    oval, oidx, opoc = og.tract_variables_batch(frames, synthetic.TV_ART_INDEX)
    assert _rel(val.cpu(), g["val64"]) <= REL
    assert torch.equal(idx.cpu(), oidx)
    assert np.array_equal(poc.cpu().numpy().astype(np.float64), g["poc64"])
    assert idx[-1, 0].tolist() == [0, 0]  # 4-way tie frame: smallest j, then smallest i


def test_tv_vs_oracle_many(asb):
    _, TV = asb
    from oracle import geometry as og
    frames = synthetic.tv_frames(3001, seed=21)
    val, idx, poc = TV.tract_variables_batch(frames.cuda(), synthetic.TV_ARTICULATORS)
    oval, oidx, opoc = og.tract_variables_batch_vectorized(frames, synthetic.TV_ART_INDEX)
    assert _rel(val.cpu(), oval) <= REL
    assert torch.equal(idx.cpu(), oidx)
    assert torch.equal(poc.cpu().double(), opoc)


def test_tv_dict_api(asb, golden_dir):
    _, TV = asb
    g = np.load(os.path.join(golden_dir, "tv.npz"))
    fr = torch.from_numpy(g["frames"])[3]
    d = {n: fr[k].T for k, n in enumerate(synthetic.TV_ARTICULATORS)}
    r = TV.calculate_vocal_tract_variables(d)
    assert r["LP"] is None and r["TTCL"] is None and r["TBCL"] is None and r["GLO"] is None
    for k, name in enumerate(("LA", "TTCD", "TBCD", "VEL")):
        assert isinstance(r[name]["value"], float)
        assert abs(r[name]["value"] - g["val64"][3, k]) <= REL * g["val64"][3, k]
        assert r[name]["poc_1"].shape == (2,)
        assert abs(r[name]["poc_1"][0].item() - g["poc64"][3, k, 0, 0]) == 0
        assert abs(r[name]["poc_2"][1].item() - g["poc64"][3, k, 1, 1]) == 0
    bad = dict(d); bad.pop("tongue")
    with pytest.raises(KeyError):
        TV.calculate_vocal_tract_variables(bad)


def test_tract_variables_driver_csv_matches_reference(golden_dir, tmp_path):
    """The frame driver (phoneme_to_articulation/__init__.py:201-297): same files, same columns in
    the same order, same ids; values within 1e-5 relative of the reference driver's own CSVs and
    the points of closest approach equal to the fp32 inputs they were picked from."""
    import pandas as pd
    from artspeech_b200 import tract_variables as TV
    g = np.load(os.path.join(golden_dir, "tv_driver.npz"))
    pred, target = torch.from_numpy(g["pred"]), torch.from_numpy(g["target"])
    frame_ids = [["%04d" % (10 + k) for k in range(6)], ["%04d" % (200 + k) for k in range(4)]]
    phonemes = [list("abcdef"), list("wxyz")]
    TV.tract_variables(["S1", "S2"], frame_ids, pred, target, [6, 4], phonemes, list(synthetic.TV_ARTICULATORS),
                       str(tmp_path))
    for sid, n in (("S1", 6), ("S2", 4)):
        got = pd.read_csv(os.path.join(str(tmp_path), sid, "tract_variables.csv"), dtype={"frame": str})
        want = pd.read_csv(os.path.join(golden_dir, f"tv_driver_{sid}.csv"), dtype={"frame": str})
        assert list(got.columns) == list(want.columns) and len(got) == len(want) == n
        for col in ("sentence", "frame", "phoneme"):
            assert list(got[col]) == list(want[col])
        for col in want.columns[3:]:
            a, b = got[col].to_numpy(dtype=np.float64), want[col].to_numpy(dtype=np.float64)
            if "_poc_" in col:
                assert np.array_equal(a.astype(np.float32), b.astype(np.float32)), col
            else:
                assert np.all(np.abs(a - b) <= 1e-5 * np.abs(b)), col


def test_config3_full_size_p2cp_and_tv(asb):
    """BASELINE configs[2]: 1 048 576 frames x 11 articulators x 50 points (4.6 GB per tensor, byte
    offsets beyond 2^32).  Size-independent properties: the frames are 256 copies of a 4096-frame
    block, so every copy must reproduce the block's result bit for bit wherever it lies in memory
    (persistent-CTA tile -> address arithmetic), the block itself is checked against the fp64 oracle,
    and P2CP of a tensor with itself is exactly zero everywhere."""
    from oracle import geometry as og
    M, TV = asb
    F, blk = 1 << 20, 4096
    base_t = synthetic.tv_frames(blk, seed=31)
    base_p = synthetic.perturbed(base_t, seed=32)
    tgt = base_t.cuda().repeat(F // blk, 1, 1, 1)
    pred = base_p.cuda().repeat(F // blk, 1, 1, 1)
    assert tgt.shape == (F, 11, 2, 50) and tgt.numel() * 4 > (1 << 32)
    p = M.p2cp_distance(pred, tgt)                                    # (F, 11)
    assert p.shape == (F, 11)
    want = og.p2cp_distance(base_p, base_t)
    assert _rel(p[:blk].cpu(), want) <= REL
    assert torch.equal(p.view(F // blk, blk, 11), p[:blk].unsqueeze(0).expand(F // blk, blk, 11))
    assert float(M.p2cp_distance(tgt, tgt).abs().max()) == 0.0
    val, idx, poc = TV.tract_variables_batch(tgt, synthetic.TV_ARTICULATORS)
    oval, oidx, opoc = og.tract_variables_batch_vectorized(base_t, synthetic.TV_ART_INDEX)
    assert _rel(val[:blk].cpu(), oval) <= REL
    assert torch.equal(idx[:blk].cpu(), oidx) and torch.equal(poc[:blk].cpu().double(), opoc)
    for t in (val, idx, poc):
        v = t.view(F // blk, blk, -1)
        assert torch.equal(v, v[:1].expand_as(v))


def test_metric_helpers_accept_cpu_tensors_and_run_on_the_gpu(asb):
    """pearsons_correlation (metrics.py:9-35, quirk of :22 included) and the EuclideanDistance criterion
    (phoneme_to_articulation/metrics.py:5-24) on CPU inputs, as the reference's evaluation loop passes
    them: computed on the CUDA device, returned on the CPU, equal to the fp64 oracle."""
    from oracle import geometry as og
    M, _ = asb
    tgt = synthetic.contours(3 * 17, 4, 50, seed=51).view(3, 17, 4, 2, 50)
    pred = synthetic.perturbed(tgt, seed=52)
    xc, yc = M.pearsons_correlation(pred, tgt)
    wx, wy = og.pearsons_correlation(pred, tgt)
    assert xc.device.type == "cpu" and xc.shape == wx.shape
    assert torch.allclose(xc.double(), wx, atol=2e-6) and torch.allclose(yc.double(), wy, atol=2e-6)
    d = M.EuclideanDistance("none")(pred, tgt)
    assert d.device.type == "cpu" and d.shape == (3, 17, 4, 50)
    want = torch.sqrt(((pred.double() - tgt.double()) ** 2).sum(dim=-2))
    assert torch.allclose(d.double(), want, rtol=1e-5, atol=1e-9)
    # the criterion's gradient (asb_point_distance_backward) against fp64 autograd of the formula
    p = pred.clone().cuda().requires_grad_(True)
    w = torch.randn(3, 17, 4, 50, generator=torch.Generator().manual_seed(3))
    (M.EuclideanDistance("none")(p, tgt.cuda()) * w.cuda()).sum().backward()
    p64 = pred.double().requires_grad_(True)
    (torch.sqrt(((p64 - tgt.double()) ** 2).sum(dim=-2)) * w.double()).sum().backward()
    assert torch.allclose(p.grad.cpu().double(), p64.grad, rtol=1e-4, atol=1e-6)
    M.EuclideanDistance("mean")(p, tgt.cuda()).backward()
    assert torch.isfinite(p.grad).all()
