"""Area-function oracle (oracle/area.py, oracle/area_c.c): pinned against the parts of the
reference's area_function.py that run without GEOS (tests/golden/area.npz), anchored by
known-answer cases for the GEOS-dependent parts, and the two restatements against each other."""
import math
import os

import numpy as np
import pytest
import torch

from artspeech_b200 import synthetic
from oracle import area as oarea
from oracle import build_c
from oracle import reference_loader as ref

ROTATED = dict(center=(0.45, 0.55), theta_rad=0.2, omega_rad=-0.15, linear_step=0.02,
               polar_step_rad=math.pi / 30)


@pytest.fixture(scope="module")
def ga(golden_dir):
    return np.load(os.path.join(golden_dir, "area.npz"))


def test_grid_matches_reference_bit_exact(ga):
    g = oarea.build_semipolar_grid(**synthetic.AREA_GRID_ARGS)
    assert g.shape == (117, 50, 2) and np.array_equal(g, ga["grid_default"])
    g2 = oarea.build_semipolar_grid(**ROTATED)
    assert np.array_equal(g2, ga["grid_rotated"])
    ends = oarea.grid_endpoints(**ROTATED)
    assert np.array_equal(ends[:, 0], g2[:, 0]) and np.array_equal(ends[:, 1], g2[:, -1])


def test_helpers_match_reference(ga):
    mids = np.array([oarea.mid_point(a, b) for a, b in ga["mid_in"]])
    assert np.array_equal(mids, ga["mid_out"])
    arg = np.array([oarea.argmatrix(ga["arg_in"], "min"), oarea.argmatrix(ga["arg_in"], "max")])
    assert np.array_equal(arg, ga["arg_out"])
    assert tuple(arg[0]) == (1, 4)  # first of the two tied minima in row-major order
    rot = np.array([oarea.rotate(np.array([0.3, -0.4]), a) for a in (0.0, 0.7, -2.1)])
    assert np.array_equal(rot, ga["rot_out"])
    with pytest.raises(AssertionError):
        oarea.argmatrix(ga["arg_in"], "median")


def test_area_function_matches_reference(ga):
    for f in range(3):
        d, fx = oarea.area_function(ga[f"int_{f}"], ga[f"ext_{f}"])
        assert np.array_equal(d, ga[f"dists_{f}"]) and np.array_equal(fx, ga[f"fx_{f}"])
        d, fx = oarea.area_function(ga[f"int_{f}"], ga[f"ext_{f}"], alpha=2.5, beta=1.5)
        assert np.array_equal(d, ga[f"dists_b_{f}"]) and np.array_equal(fx, ga[f"fx_b_{f}"])
    with pytest.raises(AssertionError):
        oarea.area_function(np.zeros((4, 2)), np.zeros((5, 2)))


@pytest.mark.skipif(not ref.available(), reason="reference checkout not present")
def test_live_reference_grid_and_area_function(ga):
    af = ref.area_function()
    g = af.build_semipolar_grid(np.array([0.5, 0.5]), 0.1, 0.05, 0.015, math.pi / 40)
    mine = oarea.build_semipolar_grid((0.5, 0.5), 0.1, 0.05, 0.015, math.pi / 40)
    assert np.array_equal(g, mine)
    d, fx = af.area_function(ga["int_1"], ga["ext_1"], alpha=1.0, beta=3.0)
    d2, fx2 = oarea.area_function(ga["int_1"], ga["ext_1"], alpha=1.0, beta=3.0)
    assert np.array_equal(d, d2) and np.array_equal(fx, fx2)


# ---------------------------------------------------------------- known answers (a10, a12)
def _default_grid():
    return oarea.build_semipolar_grid(**synthetic.AREA_GRID_ARGS)


def test_kat_straight_walls_through_mouth_lines():
    """Horizontal walls y = 0.4 / y = 0.3 over x in [0.105, 0.455] meet the vertical mouth lines
    x = 0.5 - 0.01 k, k = 5..39, at (x, 0.4) / (x, 0.3): mid-points (x, 0.35), radius 0.05."""
    grid = _default_grid()
    xs = np.linspace(0.105, 0.455, 50)
    wi = np.stack([xs, np.full(50, 0.4)], axis=1)
    we = np.stack([xs, np.full(50, 0.3)], axis=1)
    ii, ee, lines = oarea.intersect_semipolar_grid(wi, we, grid, return_lines=True)
    ks = np.arange(5, 40)
    assert np.array_equal(lines, 67 + ks)      # mouth lines follow 50 larynx + 17 polar lines
    want_x = grid[67 + ks, 0, 0]
    np.testing.assert_allclose(ii[:, 0], want_x, rtol=0, atol=1e-15)
    np.testing.assert_allclose(ee[:, 0], want_x, rtol=0, atol=1e-15)
    np.testing.assert_allclose(ii[:, 1], 0.4, rtol=0, atol=1e-15)
    np.testing.assert_allclose(ee[:, 1], 0.3, rtol=0, atol=1e-15)
    d, fx = oarea.area_function(ii, ee)
    np.testing.assert_allclose(fx, math.pi * 0.05 ** 2, rtol=1e-13)
    np.testing.assert_allclose(d, 0.01 * np.arange(35), rtol=0, atol=1e-13)
    xfx = oarea.evenly_spaced_fx(d, fx)
    assert xfx.dtype == torch.float32 and tuple(xfx.shape) == (2, 200)
    np.testing.assert_allclose(xfx[0].numpy(), np.linspace(0, 0.34, 200), rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(xfx[1].numpy(), math.pi * 0.0025, rtol=1e-6)


def test_kat_closest_of_several_crossings_and_vertex_touch():
    grid = _default_grid()
    line = 67 + 20                                  # the mouth line x = 0.3, y from 0.5 to 0.1
    x0 = grid[line, 0, 0]
    # internal wall: a zig-zag crossing x = x0 at y = 0.45, 0.35 and 0.25
    wi = np.array([[x0 - 0.004, 0.45], [x0 + 0.004, 0.45], [x0 + 0.004, 0.35], [x0 - 0.004, 0.35],
                   [x0 - 0.004, 0.25], [x0 + 0.004, 0.25]])
    # external wall: one vertex exactly on the line at y = 0.33 (a touch shared by two segments)
    we = np.array([[x0 - 0.003, 0.31], [x0, 0.33], [x0 - 0.003, 0.34]])
    ii, ee, lines = oarea.intersect_semipolar_grid(wi, we, grid, return_lines=True)
    assert list(lines) == [line]
    assert ee[0, 0] == x0 and ee[0, 1] == 0.33               # exact vertex coordinates, once
    np.testing.assert_allclose(ii[0], [x0, 0.35], rtol=0, atol=1e-15)   # closest of the three
    pts = oarea.polyline_intersections(grid[line][[0, -1]], we)
    assert pts == [(x0, 0.33)]


def test_kat_single_contact_uses_external_wall_ends_for_both_walls():
    """area_function.py:199-221: with one wall missed, the other wall's intersection is paired
    with coords[-j_min] of the missed wall, j_min chosen by distance to the EXTERNAL wall's ends."""
    grid = _default_grid()
    line = 67 + 10                                  # x = 0.4
    x0 = grid[line, 0, 0]
    wi = np.array([[x0 - 0.004, 0.40], [x0 + 0.004, 0.40]])       # crosses at (x0, 0.40)
    we = np.array([[x0 + 0.002, 0.10 - 0.05], [x0 + 0.004, 0.38]])  # misses every grid line
    ii, ee = oarea.intersect_semipolar_grid(wi, we, grid)
    assert len(ii) == 1
    np.testing.assert_allclose(ii[0], [x0, 0.40], atol=1e-15)
    assert np.array_equal(ee[0], we[-1])            # nearer end -> j_min = 1 -> coords[-1]
    # external contact only: distances are still taken to the external wall's own ends
    we2 = np.array([[x0 - 0.004, 0.30], [x0 + 0.012, 0.30]])      # crosses x0 near its first end
    wi2 = np.array([[0.95, 0.05], [0.96, 0.05]])                  # misses the grid
    ii2, ee2, lines = oarea.intersect_semipolar_grid(wi2, we2, grid, return_lines=True)
    k = list(lines).index(line)
    np.testing.assert_allclose(ee2[k], [x0, 0.30], atol=1e-15)
    assert np.array_equal(ii2[k], wi2[0])           # j_min = 0 -> internal coords[0]


def test_collinear_overlap_contributes_nothing():
    grid = _default_grid()
    line = 67 + 10
    x0 = grid[line, 0, 0]
    wi = np.array([[x0, 0.45], [x0, 0.35]])         # lies on the grid line
    we = np.array([[0.95, 0.05], [0.96, 0.05]])
    ii, ee = oarea.intersect_semipolar_grid(wi, we, grid)
    assert ii.shape == (0, 2) and ee.shape == (0, 2)


def test_evenly_spaced_fx_is_linear_interpolation():
    rng = np.random.default_rng(0)
    x = np.cumsum(np.concatenate([[0.0], rng.random(30) + 0.01]))
    fx = rng.random(31)
    got = oarea.evenly_spaced_fx(x, fx, 57, dtype=torch.float64).numpy()
    xs = np.linspace(x[0], x[-1], 57)
    assert np.array_equal(got[0], xs)
    np.testing.assert_allclose(got[1], np.interp(xs, x, fx), rtol=1e-13, atol=1e-15)
    assert got[1, 0] == fx[0] and got[1, -1] == fx[-1]


# ---------------------------------------------------------------- C restatement vs NumPy restatement
def test_c_oracle_equals_numpy_oracle_on_synthetic_frames():
    grid = _default_grid()
    ends = oarea.grid_endpoints(**synthetic.AREA_GRID_ARGS)
    wi, we = synthetic.walls(24, seed=21)
    r = build_c.area_function_batch(wi.numpy(), we.numpy(), ends)
    single = 0
    for f in range(24):
        a, b = wi[f].numpy().astype(np.float64), we[f].numpy().astype(np.float64)
        xfx, K, ii, ee, lines = oarea.vocal_tract_area_function(a, b, grid)
        assert K == r["K"][f] and np.array_equal(lines, r["lines"][f, :K])
        assert np.all(r["lines"][f, K:] == -1)
        assert np.array_equal(ii, r["pts"][f, :K, 0]) and np.array_equal(ee, r["pts"][f, :K, 1])
        np.testing.assert_allclose(r["xfx"][f], xfx, rtol=1e-14, atol=1e-18)
        if f < 4:   # every sub-segment of the 50-sample grid lines gives the same points
            ii2, ee2 = oarea.intersect_semipolar_grid(a, b, grid, polyline=True)
            np.testing.assert_allclose(ii2, ii, rtol=0, atol=1e-15)
            np.testing.assert_allclose(ee2, ee, rtol=0, atol=1e-15)
        # single-contact lines: an intersection point that is a wall end
        single += int(np.any(np.all(ee == b[0], axis=1) | np.all(ee == b[-1], axis=1)
                             | np.all(ii == a[0], axis=1) | np.all(ii == a[-1], axis=1)))
    assert single > 0, "the synthetic walls should exercise the single-contact branch"


def test_c_oracle_threads_and_degenerate_frames():
    ends = oarea.grid_endpoints(**synthetic.AREA_GRID_ARGS)
    wi, we = synthetic.walls(64, seed=5)
    wi, we = wi.numpy().copy(), we.numpy().copy()
    wi[3] = 0.97; we[3] = 0.98                       # misses every line -> K = 0
    r1 = build_c.area_function_batch(wi, we, ends, threads=1)
    r4 = build_c.area_function_batch(wi, we, ends, threads=4)
    for k in r1:
        assert np.array_equal(r1[k], r4[k], equal_nan=True)
    assert r1["K"][3] == 0 and np.all(np.isnan(r1["xfx"][3]))
    assert r1["K"].max() <= 117 and np.all(np.isfinite(r1["xfx"][np.arange(64) != 3]))


def test_product_host_helpers_match_reference(ga):
    """The package's host-side grid builder / helpers (configuration, no CUDA involved)."""
    from artspeech_b200 import area_function as af
    assert np.array_equal(af.build_semipolar_grid(**synthetic.AREA_GRID_ARGS), ga["grid_default"])
    assert np.array_equal(af.build_semipolar_grid(**ROTATED), ga["grid_rotated"])
    assert np.array_equal(np.array([af.mid_point(a, b) for a, b in ga["mid_in"]]), ga["mid_out"])
    assert np.array_equal(np.array([af.argmatrix(ga["arg_in"], "min"), af.argmatrix(ga["arg_in"], "max")]),
                          ga["arg_out"])
    assert np.array_equal(np.array([af.rotate(np.array([0.3, -0.4]), a) for a in (0.0, 0.7, -2.1)]), ga["rot_out"])
    ends = af.grid_endpoints(ga["grid_rotated"])
    assert np.array_equal(ends, oarea.grid_endpoints(**ROTATED))
