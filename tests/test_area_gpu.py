"""Area-function CUDA path (asb_area_function and the two split entry points, through the C ABI)
against the fp64 oracle: the touched grid lines K / their indices / the kept intersection points
bit-exact, the resampled profile within 1e-5 relative (BASELINE.json north_star)."""
import math
import os

import numpy as np
import pytest
import torch

from artspeech_b200 import area_function as af
from artspeech_b200 import synthetic
from oracle import area as oarea
from oracle import build_c

pytestmark = pytest.mark.gpu

RTOL = 1e-5   # north_star: "P2CP, TVs and area function within 1e-5 relative"


@pytest.fixture(scope="module")
def grid():
    return af.build_semipolar_grid(**synthetic.AREA_GRID_ARGS)


@pytest.fixture(scope="module")
def dgrid(grid):
    return af.DeviceGrid(grid, torch.device("cuda", 0))


def _check_against_c_oracle(wi, we, grid, dgrid, **kw):
    ends = af.grid_endpoints(grid)
    want = build_c.area_function_batch(wi.numpy(), we.numpy(), ends, threads=os.cpu_count() or 1,
                                       alpha=kw.get("alpha", np.pi), beta=kw.get("beta", 2.0),
                                       n_samples=kw.get("n_samples", 200))
    got = af.area_function_batch(wi.cuda(), we.cuda(), dgrid, return_points=True, **kw)
    torch.cuda.synchronize()
    K = got["K"].cpu().numpy()
    assert np.array_equal(K, want["K"]), "touched grid lines differ"
    assert np.array_equal(got["lines"].cpu().numpy(), want["lines"])
    assert np.array_equal(got["int_pts"].cpu().numpy(), want["pts"][:, :, 0])      # bit-exact fp64
    assert np.array_equal(got["ext_pts"].cpu().numpy(), want["pts"][:, :, 1])
    xfx = got["xfx"].cpu().numpy().astype(np.float64)
    ok = K >= 2
    assert np.all(np.isnan(xfx[~ok])) and np.all(np.isnan(want["xfx"][~ok]))
    a, b = xfx[ok], want["xfx"][ok]
    assert np.all(np.isfinite(a)) and np.all(np.isfinite(b))
    rel = np.abs(a - b) / np.maximum(np.abs(b), 1e-30)
    rel[b == 0] = np.abs(a[b == 0])           # the first abscissa is exactly 0
    assert rel.max() <= RTOL, f"profile off by {rel.max():.2e} relative"
    # without the optional outputs the profile is the same
    lean = af.area_function_batch(wi.cuda(), we.cuda(), dgrid, **kw)
    assert torch.equal(lean["K"], got["K"])
    assert torch.equal(torch.nan_to_num(lean["xfx"]), torch.nan_to_num(got["xfx"]))
    return rel.max(), K


def test_synthetic_frames_match_oracle(grid, dgrid):
    wi, we = synthetic.walls(4096, seed=31)
    rel, K = _check_against_c_oracle(wi, we, grid, dgrid)
    assert 90 <= K.min() and K.max() <= 117
    print(f"area function: 4096 frames, K {K.min()}..{K.max()}, max rel {rel:.2e}")


def test_other_parameters_and_rotated_grid():
    kw = dict(center=(0.45, 0.55), theta_rad=0.2, omega_rad=-0.15, linear_step=0.02, polar_step_rad=math.pi / 30)
    grid = af.build_semipolar_grid(**kw)
    dgrid = af.DeviceGrid(grid, torch.device("cuda", 0))
    wi, we = synthetic.walls(512, n_points=37, seed=32)
    _check_against_c_oracle(wi, we, grid, dgrid, alpha=2.5, beta=1.5, n_samples=77)
    wi, we = synthetic.walls(64, n_points=64, seed=33)
    _check_against_c_oracle(wi, we, grid, dgrid, n_samples=2)


def test_lattice_walls_exercise_exact_predicates(grid, dgrid):
    """Vertices snapped to multiples of 1/64 land exactly on grid lines (x = 0.25, y = 0.75, the
    centre shared by 19 lines), on each other and on collinear runs: zero orientations, touches,
    duplicate points and distance ties must all be decided as the fp64 oracle decides them."""
    g = torch.Generator().manual_seed(7)
    F = 2048
    wi = torch.randint(6, 60, (F, 50, 2), generator=g).float() / 64
    we = torch.randint(6, 60, (F, 50, 2), generator=g).float() / 64
    wi[::3, 7] = 0.5            # a vertex on the grid centre
    we[::5, 11, 0] = 0.25
    wi[::7, 20:24, 1] = 0.75    # a run lying on a larynx line
    _check_against_c_oracle(wi, we, grid, dgrid)
    # smooth walls snapped to the lattice: realistic shapes with many exact touches
    a, b = synthetic.walls(1024, seed=8)
    snap = lambda t: torch.round(t * 128) / 128
    _check_against_c_oracle(snap(a), snap(b), grid, dgrid)


def test_degenerate_frames(grid, dgrid):
    wi, we = synthetic.walls(8, seed=9)
    wi, we = wi.clone(), we.clone()
    wi[0] = 0.97; we[0] = 0.98                 # touches nothing -> K = 0, NaN profile
    wi[1, :] = wi[1, :1]; we[1, :] = we[1, :1]  # collapsed to one point each
    we[2] = 0.98                                # only the internal wall is met on every line
    wi[3] = 0.98                                # only the external wall is met
    _, K = _check_against_c_oracle(wi, we, grid, dgrid)
    assert K[0] == 0 and K[2] > 50 and K[3] > 50
    empty = af.area_function_batch(wi[:0].cuda(), we[:0].cuda(), dgrid)
    assert tuple(empty["xfx"].shape) == (0, 2, 200)
    with pytest.raises(ValueError):
        af.area_function_batch(torch.zeros(2, 65, 2).cuda(), torch.zeros(2, 65, 2).cuda(), dgrid)
    with pytest.raises(ValueError):
        af.area_function_batch(torch.zeros(2, 50, 2), torch.zeros(2, 49, 2), dgrid)


def test_reference_signature_functions(grid, golden_dir):
    ga = np.load(os.path.join(golden_dir, "area.npz"))
    assert np.array_equal(grid, ga["grid_default"])           # product grid == reference grid, bit-exact
    for f in range(3):
        wi, we = ga["wall_int"][f], ga["wall_ext"][f]
        ii, ee = af.intersect_semipolar_grid(wi, we, grid)
        assert ii.dtype == np.float64 and np.array_equal(ii, ga[f"int_{f}"]) and np.array_equal(ee, ga[f"ext_{f}"])
        # area_function(): golden values are outputs of the reference's own function
        d, fx = af.area_function(ii, ee)
        np.testing.assert_allclose(d, ga[f"dists_{f}"], rtol=1e-13, atol=0)
        np.testing.assert_allclose(fx, ga[f"fx_{f}"], rtol=1e-13, atol=0)
        d2, fx2 = af.area_function(ii, ee, alpha=2.5, beta=1.5)
        np.testing.assert_allclose(d2, ga[f"dists_b_{f}"], rtol=1e-13)
        np.testing.assert_allclose(fx2, ga[f"fx_b_{f}"], rtol=1e-12)
        xfx = af.evenly_spaced_fx(d, fx)
        want = oarea.evenly_spaced_fx(ga[f"dists_{f}"], ga[f"fx_{f}"])
        assert xfx.dtype == torch.float32 and tuple(xfx.shape) == (2, 200) and not xfx.is_cuda
        np.testing.assert_allclose(xfx.numpy(), want.numpy(), rtol=2e-7, atol=0)
    with pytest.raises(AssertionError):
        af.area_function(np.zeros((4, 2)), np.zeros((5, 2)))
    # long polylines through the split kernels (K > 32 chunks, K not a multiple of 32)
    rng = np.random.default_rng(1)
    p, q = rng.random((333, 2)), rng.random((333, 2))
    d, fx = af.area_function(p, q)
    dw, fw = oarea.area_function(p, q)
    np.testing.assert_allclose(d, dw, rtol=1e-13)
    np.testing.assert_allclose(fx, fw, rtol=1e-13)
    got = af.evenly_spaced_fx(dw, fw, 1000).numpy()
    np.testing.assert_allclose(got[1], np.interp(np.linspace(dw[0], dw[-1], 1000), dw, fw), rtol=2e-7)


def test_known_answer_straight_walls(grid, dgrid):
    xs = np.linspace(0.105, 0.455, 50)
    wi = torch.tensor(np.stack([xs, np.full(50, 0.4)], axis=1), dtype=torch.float32)
    we = torch.tensor(np.stack([xs, np.full(50, 0.3)], axis=1), dtype=torch.float32)
    r = af.area_function_batch(wi, we, dgrid, return_points=True)
    K = int(r["K"])
    assert K == 35 and r["lines"][:K].tolist() == list(range(72, 107))
    xfx = r["xfx"].cpu().double().numpy()
    radius = (np.float32(0.4).astype(np.float64) - np.float32(0.3).astype(np.float64)) / 2
    np.testing.assert_allclose(xfx[1], math.pi * radius ** 2, rtol=RTOL)
    np.testing.assert_allclose(xfx[0], np.linspace(0, 0.34, 200), rtol=RTOL, atol=1e-9)


def test_full_size_properties(dgrid, grid):
    """BASELINE config 3 size (1M frames): determinism, structure of the output, and the C oracle
    on a strided subsample."""
    F = 1 << 20
    base_i, base_e = synthetic.walls(1 << 14, seed=41)
    reps = F // base_i.shape[0]
    wi = base_i.cuda().repeat(reps, 1, 1)
    we = base_e.cuda().repeat(reps, 1, 1)
    # make the copies distinct: shift each repetition by a small multiple of one fp32-exact step
    shift = (torch.arange(reps, device="cuda", dtype=torch.float32) * (2.0 ** -12)).repeat_interleave(base_i.shape[0])
    wi[:, :, 1] -= shift[:, None] * 0.5
    we[:, :, 1] -= shift[:, None] * 0.5
    r1 = af.area_function_batch(wi, we, dgrid)
    r2 = af.area_function_batch(wi, we, dgrid)
    torch.cuda.synchronize()
    assert torch.equal(r1["K"], r2["K"]) and torch.equal(torch.nan_to_num(r1["xfx"]), torch.nan_to_num(r2["xfx"]))
    K, xfx = r1["K"], r1["xfx"]
    ok = K >= 2
    assert int(K.max()) <= 117 and ok.float().mean() > 0.99
    x = xfx[ok][:, 0]
    assert torch.all(x[:, 0] == 0) and torch.all(x[:, 1:] > x[:, :-1])         # a strictly increasing linspace
    step = x[:, -1:] / 199
    assert torch.all((x - step * torch.arange(200, device="cuda")).abs() <= 4e-7 * x[:, -1:] + 1e-12)
    assert torch.all(xfx[ok][:, 1] >= 0)
    sel = torch.arange(0, F, 257)
    ends = af.grid_endpoints(grid)
    want = build_c.area_function_batch(wi[sel].cpu().numpy(), we[sel].cpu().numpy(), ends,
                                       threads=os.cpu_count() or 1, want_points=False)
    assert np.array_equal(K[sel].cpu().numpy(), want["K"])
    a, b = xfx[sel].cpu().double().numpy(), want["xfx"]
    m = np.isfinite(b)
    assert np.array_equal(m, np.isfinite(a))
    rel = np.abs(a[m] - b[m]) / np.maximum(np.abs(b[m]), 1e-30)
    rel[b[m] == 0] = 0
    assert rel.max() <= RTOL
