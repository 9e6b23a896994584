"""Exact-arithmetic anchor for rows a10 / a12 (shapely/GEOS is absent, so the fp64
restatement in oracle/area.py cannot be pinned against the reference's own outputs).

GEOS decides orientations with an EXACT predicate; the restatement uses the plain fp64 expression.
Here every orientation sign, every touch / crossing / miss classification and every selected
intersection of the restatement is re-derived in exact rational arithmetic (Python integers scaled
to a common binary exponent for the signs, ``fractions.Fraction`` for the crossing points) on the
same inputs the GPU parity tests use - lattice walls that land exactly on grid lines, on the grid
centre and on collinear runs, lattice-snapped smooth walls, and plain synthetic walls - so the
fp64 decisions are verified by code that shares nothing with oracle/area.py but the input arrays.
"""
import math
from fractions import Fraction

import numpy as np
import torch

from artspeech_b200 import synthetic
from oracle import area as oarea


def _as_scaled_ints(values):
    """Exact integers n_k with values[k] = n_k * 2**e for one common exponent e."""
    parts = [math.frexp(float(v)) for v in values]            # v = m * 2**x, 0.5 <= |m| < 1
    ints = [(int(m * (1 << 53)), x - 53) for m, x in parts]   # exact: m has 53 significant bits
    e = min(x for _, x in ints)
    return [n << (x - e) for n, x in ints], e


def exact_orient_sign(a, b, c):
    """Sign of (b - a) x (c - a) in exact integer arithmetic."""
    (ax, ay, bx, by, cx, cy), _ = _as_scaled_ints([a[0], a[1], b[0], b[1], c[0], c[1]])
    det = (bx - ax) * (cy - ay) - (by - ay) * (cx - ax)
    return (det > 0) - (det < 0)


def exact_segment_intersection(p0, p1, q0, q1):
    """('none' | 'touch' | 'cross', point) with the point as a pair of Fractions; same case order as
    the published behaviour restated in oracle.area.segment_intersection (collinear overlap = none)."""
    s1, s2 = exact_orient_sign(p0, p1, q0), exact_orient_sign(p0, p1, q1)
    if s1 * s2 > 0:
        return "none", None
    s3, s4 = exact_orient_sign(q0, q1, p0), exact_orient_sign(q0, q1, p1)
    if s3 * s4 > 0:
        return "none", None
    if s1 == 0 and s2 == 0 and s3 == 0 and s4 == 0:
        return "none", None
    F = lambda p: (Fraction(float(p[0])), Fraction(float(p[1])))
    for s, p in ((s1, q0), (s2, q1), (s3, p0), (s4, p1)):
        if s == 0:
            return "touch", F(p)
    P0, P1, Q0, Q1 = F(p0), F(p1), F(q0), F(q1)
    d1 = (P1[0] - P0[0]) * (Q0[1] - P0[1]) - (P1[1] - P0[1]) * (Q0[0] - P0[0])
    d2 = (P1[0] - P0[0]) * (Q1[1] - P0[1]) - (P1[1] - P0[1]) * (Q1[0] - P0[0])
    t = d1 / (d1 - d2)
    return "cross", (Q0[0] + t * (Q1[0] - Q0[0]), Q0[1] + t * (Q1[1] - Q0[1]))


def _check_frame(wall_int, wall_ext, ends):
    """Every (grid line, wall segment) pair of one frame: fp64 restatement vs exact arithmetic."""
    counts = {"none": 0, "touch": 0, "cross": 0}
    worst = 0.0
    for wall in (wall_int, wall_ext):
        for line in ends:
            p0, p1 = line[0], line[1]
            for k in range(len(wall) - 1):
                got = oarea.segment_intersection(p0, p1, wall[k], wall[k + 1])
                kind, pt = exact_segment_intersection(p0, p1, wall[k], wall[k + 1])
                counts[kind] += 1
                if kind == "none":
                    assert got is None, (p0, p1, wall[k], wall[k + 1], got)
                    continue
                assert got is not None, (kind, p0, p1, wall[k], wall[k + 1])
                if kind == "touch":      # the endpoint's own coordinates, bit for bit
                    assert Fraction(got[0]) == pt[0] and Fraction(got[1]) == pt[1]
                else:
                    err = max(abs(Fraction(got[0]) - pt[0]), abs(Fraction(got[1]) - pt[1]))
                    worst = max(worst, float(err))
    return counts, worst


def _selected_points_exact(wall_int, wall_ext, ends):
    """intersect_semipolar_grid (area_function.py:175-223) re-derived with exact intersections and
    exact squared distances: the chosen (line, internal point, external point) triples."""
    def hits(line, wall):
        out = []
        for k in range(len(wall) - 1):
            kind, pt = exact_segment_intersection(line[0], line[1], wall[k], wall[k + 1])
            if kind != "none" and pt not in out:
                out.append(pt)
        return out

    F = lambda p: (Fraction(float(p[0])), Fraction(float(p[1])))
    d2 = lambda a, b: (a[0] - b[0]) ** 2 + (a[1] - b[1]) ** 2
    default = [F(wall_ext[0]), F(wall_ext[-1])]
    chosen = []
    for l, line in enumerate(ends):
        li, le = hits(line, wall_int), hits(line, wall_ext)
        if not li and not le:
            continue
        pi = pe = None
        if li:
            cmp_to = le if le else default
            best = min(((d2(a, b), i, j) for i, a in enumerate(li) for j, b in enumerate(cmp_to)))
            pi = li[best[1]]
            if not le:
                pe = F(wall_ext[-best[2]])
        if le:
            cmp_to = li if li else default
            best = min(((d2(a, b), i, j) for i, a in enumerate(le) for j, b in enumerate(cmp_to)))
            pe = le[best[1]]
            if not li:
                pi = F(wall_int[-best[2]])
        chosen.append((l, pi, pe))
    return chosen


def _lattice_frames():
    g = torch.Generator().manual_seed(7)   # the frames of test_area_gpu.test_lattice_walls_...
    F = 2048
    wi = torch.randint(6, 60, (F, 50, 2), generator=g).float() / 64
    we = torch.randint(6, 60, (F, 50, 2), generator=g).float() / 64
    wi[::3, 7] = 0.5
    we[::5, 11, 0] = 0.25
    wi[::7, 20:24, 1] = 0.75
    return wi.double().numpy(), we.double().numpy()


def test_fp64_decisions_equal_exact_arithmetic_on_lattice_walls():
    ends = oarea.grid_endpoints(**synthetic.AREA_GRID_ARGS)
    wi, we = _lattice_frames()
    total = {"none": 0, "touch": 0, "cross": 0}
    for f in (0, 15, 21, 105):            # 0: centre vertex + x = 0.25 + collinear run; 15, 21, 105: two of three
        counts, worst = _check_frame(wi[f], we[f], ends)
        for k in total:
            total[k] += counts[k]
        assert worst <= 1e-15, worst
    assert total["touch"] >= 40 and total["cross"] > 1000, total   # the special cases are really there


def test_fp64_decisions_equal_exact_arithmetic_on_smooth_walls():
    ends = oarea.grid_endpoints(**synthetic.AREA_GRID_ARGS)
    a, b = synthetic.walls(3, seed=8)
    snap = lambda t: (torch.round(t * 128) / 128).double().numpy()
    for wi, we in ((snap(a), snap(b)), (a.double().numpy(), b.double().numpy())):
        for f in range(2):
            counts, worst = _check_frame(wi[f], we[f], ends)
            assert counts["cross"] > 100 and worst <= 1e-15, (counts, worst)


def test_selected_pairs_equal_exact_selection():
    """The K touched lines and the kept intersection per wall (distance argmin with the reference's
    first-occurrence tie-break) as exact rational arithmetic selects them."""
    grid = oarea.build_semipolar_grid(**synthetic.AREA_GRID_ARGS)
    ends = oarea.grid_endpoints(**synthetic.AREA_GRID_ARGS)
    lw_i, lw_e = _lattice_frames()
    a, b = synthetic.walls(2, seed=21)
    cases = [(lw_i[0], lw_e[0]), (lw_i[21], lw_e[21]), (a[0].double().numpy(), b[0].double().numpy()),
             (a[1].double().numpy(), b[1].double().numpy())]
    for wi, we in cases:
        ii, ee, lines = oarea.intersect_semipolar_grid(wi, we, grid, return_lines=True)
        chosen = _selected_points_exact(wi, we, ends)
        assert [c[0] for c in chosen] == lines.tolist()
        for k, (_, pi, pe) in enumerate(chosen):
            assert abs(Fraction(float(ii[k, 0])) - pi[0]) <= Fraction(1, 10 ** 15)
            assert abs(Fraction(float(ii[k, 1])) - pi[1]) <= Fraction(1, 10 ** 15)
            assert abs(Fraction(float(ee[k, 0])) - pe[0]) <= Fraction(1, 10 ** 15)
            assert abs(Fraction(float(ee[k, 1])) - pe[1]) <= Fraction(1, 10 ** 15)


def test_evenly_spaced_fx_equals_exact_interpolation():
    """a12 in rational arithmetic: the vertical through xs meets the polyline (x, fx) at the linear
    interpolant of the first segment containing it."""
    rng = np.random.default_rng(4)
    x = np.cumsum(rng.random(40) * 0.02)
    fx = rng.random(40)
    got = oarea.evenly_spaced_fx(x, fx, n_samples=64, dtype=torch.float64).numpy()
    X, FX = [Fraction(float(v)) for v in x], [Fraction(float(v)) for v in fx]
    for k in range(64):
        xv = Fraction(float(got[0, k]))
        s = next(s for s in range(39) if X[s] <= xv <= X[s + 1])
        want = FX[s] + (xv - X[s]) / (X[s + 1] - X[s]) * (FX[s + 1] - FX[s])
        assert abs(Fraction(float(got[1, k])) - want) <= Fraction(1, 10 ** 14), k
    assert got[0, 0] == x[0] and got[0, -1] == x[-1]
