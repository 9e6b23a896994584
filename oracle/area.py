"""NumPy fp64 restatement of the reference's ``area_function.py`` (rows a9-a12 of SURVEY.md 8a).

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

PARITY STATUS
  * ``rotate`` / ``build_semipolar_grid`` / ``mid_point`` / ``argmatrix`` / ``area_function``:
    PINNED - the reference's own functions run in the build container once ``funcy``,
    ``shapely`` and ``vt_tools`` are stubbed and ``np.float`` is aliased
    (``oracle/reference_loader.area_function``); their outputs are in
    ``tests/golden/area.npz`` (``oracle/make_golden.py``).
  * ``intersect_semipolar_grid`` / ``evenly_spaced_fx``: PARITY UNPINNED.  Their arithmetic lives in
    shapely==2.0.1 / GEOS (requirements.txt:17) and ``vt_tools.metrics`` (un-vendored, unpinned,
    requirements.txt:37), neither of which is in /root/reference or installable here, the module
    has no caller, no test and no fixture in the reference.  What is restated is the published
    behaviour of those calls: ``LineString.intersects/intersection`` = all points shared by two
    polylines (proper crossings and touches), ``vt_tools.metrics.euclidean`` = sqrt(dx^2+dy^2),
    ``distance_matrix`` = all pairwise Euclidean distances.  Known-answer tests with analytically
    known intersections (tests/test_oracle_area.py) anchor the restatement instead.

DEFINITIONS where the reference itself is undefined (it raises, or GEOS leaves it open):
  * a grid line is the segment from its first to its last sample (the 50 ``linspace`` samples are
    collinear to 1 ulp; ``polyline=True`` intersects every sub-segment instead, the two agree on
    the synthetic inputs and this is tested);
  * orientation predicates are the plain fp64 expression ``orient`` below (GEOS uses an exact
    predicate; they differ only when |det| < ~1e-17);
  * the intersections of one grid line with one wall are ordered by wall-segment index (GEOS does
    not specify the order of a MultiPoint; it only matters for exact distance ties);
  * a wall segment collinear with the grid line (GEOS would return a LineString and the reference
    would raise AttributeError at area_function.py:191/197) contributes nothing;
  * ``evenly_spaced_fx`` with repeated abscissae (the reference raises) takes the first segment
    that contains the sample.
"""
from __future__ import annotations

import numpy as np
import torch


# --------------------------------------------------------------------------- a9: the grid
def rotate(point, ang_rad):
    """area_function.py:12-28 - multiply by [[cos, sin], [-sin, cos]] (a clockwise rotation)."""
    c, s = np.cos(ang_rad), np.sin(ang_rad)
    return np.array([[c, s], [-s, c]]) @ np.asarray(point, dtype=np.float64)


def grid_endpoints(center, theta_rad, omega_rad, linear_step, polar_step_rad):
    """Internal/external end of every grid line, (n_lines, 2, 2) float64, in the order of
    area_function.py:83-108: larynx lines reversed, polar lines reversed, mouth lines."""
    center = np.asarray(center, dtype=np.float64)
    # mouth cavity, :44-55: x = 0, -step, ... > -0.5, internal y = 0, external y = -0.4
    xs = np.arange(0., -0.5, -linear_step)
    mouth_int = np.array([rotate((x, 0.), theta_rad) + center for x in xs])
    mouth_ext = np.array([rotate((x, -0.4), theta_rad) + center for x in xs])
    # larynx cavity, :57-68: y = 0, step, ... < 0.5, internal x = 0, external x = 0.4
    ys = np.arange(0., 0.5, linear_step)
    larynx_int = np.array([rotate((0., y), omega_rad) + center for y in ys])
    larynx_ext = np.array([rotate((0.4, y), omega_rad) + center for y in ys])
    # polar part, :70-79
    angles = np.arange(theta_rad - polar_step_rad, -(np.pi / 2) + omega_rad, -polar_step_rad)
    polar_ext = np.array([rotate((0., -0.4), a) + center for a in angles]).reshape(-1, 2)
    polar_int = np.zeros((len(polar_ext), 2)) + center
    ends_int = np.concatenate([larynx_int[::-1], polar_int[::-1], mouth_int], axis=0)
    ends_ext = np.concatenate([larynx_ext[::-1], polar_ext[::-1], mouth_ext], axis=0)
    return np.stack([ends_int, ends_ext], axis=1)


def build_semipolar_grid(center, theta_rad, omega_rad, linear_step, polar_step_rad, grid_res=50):
    """area_function.py:31-110 -> (n_lines, grid_res, 2) float64."""
    ends = grid_endpoints(center, theta_rad, omega_rad, linear_step, polar_step_rad)
    lines = []
    for (x_int, y_int), (x_ext, y_ext) in ends:
        lines.append(np.array([np.linspace(x_int, x_ext, grid_res),
                               np.linspace(y_int, y_ext, grid_res)]).T)
    return np.array(lines)


# --------------------------------------------------------------------------- third-party pieces
def euclidean(u, v):
    """vt_tools.metrics.euclidean (call sites area_function.py:128,139)."""
    return np.sqrt((u[0] - v[0]) ** 2 + (u[1] - v[1]) ** 2)


def distance_matrix(A, B):
    """vt_tools.metrics.distance_matrix (call sites :207,216): (len(A), len(B)) Euclidean."""
    A = np.asarray(A, dtype=np.float64).reshape(-1, 2)
    B = np.asarray(B, dtype=np.float64).reshape(-1, 2)
    dx = A[:, None, 0] - B[None, :, 0]
    dy = A[:, None, 1] - B[None, :, 1]
    return np.sqrt(dx * dx + dy * dy)


def orient(ax, ay, bx, by, cx, cy):
    """Signed area predicate, plain fp64 (no fused multiply-add): > 0 if c is left of a->b."""
    return (bx - ax) * (cy - ay) - (by - ay) * (cx - ax)


def segment_intersection(p0, p1, q0, q1):
    """The point shared by segments p0p1 and q0q1, or None (disjoint, or collinear overlap).
    Touching at an endpoint returns that endpoint's exact coordinates; a proper crossing is
    q0 + t (q1 - q0) with t = d1 / (d1 - d2), d1/d2 the signed areas of q0/q1 about p0p1."""
    d1 = orient(p0[0], p0[1], p1[0], p1[1], q0[0], q0[1])
    d2 = orient(p0[0], p0[1], p1[0], p1[1], q1[0], q1[1])
    if (d1 > 0 and d2 > 0) or (d1 < 0 and d2 < 0):
        return None
    d3 = orient(q0[0], q0[1], q1[0], q1[1], p0[0], p0[1])
    d4 = orient(q0[0], q0[1], q1[0], q1[1], p1[0], p1[1])
    if (d3 > 0 and d4 > 0) or (d3 < 0 and d4 < 0):
        return None
    if d1 == 0 and d2 == 0 and d3 == 0 and d4 == 0:
        return None
    if d1 == 0:
        return (float(q0[0]), float(q0[1]))
    if d2 == 0:
        return (float(q1[0]), float(q1[1]))
    if d3 == 0:
        return (float(p0[0]), float(p0[1]))
    if d4 == 0:
        return (float(p1[0]), float(p1[1]))
    t = d1 / (d1 - d2)
    return (float(q0[0] + t * (q1[0] - q0[0])), float(q0[1] + t * (q1[1] - q0[1])))


def polyline_intersections(line, wall):
    """All points shared by two polylines, ordered by (wall segment, line segment), exact
    duplicates removed: what ``LineString(line).intersection(LineString(wall))`` holds."""
    pts = []
    for k in range(len(wall) - 1):
        for s in range(len(line) - 1):
            p = segment_intersection(line[s], line[s + 1], wall[k], wall[k + 1])
            if p is not None and p not in pts:
                pts.append(p)
    return pts


# --------------------------------------------------------------------------- a10: intersections
def argmatrix(mtx, minmax):
    """area_function.py:162-172 - (row, column) of the first extreme element in row-major order."""
    assert minmax in ["min", "max"]
    flat = np.argmin(mtx) if minmax == "min" else np.argmax(mtx)
    _, m = mtx.shape
    return flat // m, flat % m


def intersect_semipolar_grid(internal_wall, external_wall, semipolar_grid, polyline=False,
                             return_lines=False):
    """area_function.py:175-223.  walls (Nw, 2); grid (n_lines, grid_res, 2).
    Returns (internal_intersec (K,2), external_intersec (K,2)) float64, one pair per grid line
    that touches at least one wall, in grid order; with ``return_lines`` also the line indices."""
    internal_wall = np.asarray(internal_wall, dtype=np.float64)
    external_wall = np.asarray(external_wall, dtype=np.float64)
    grid = np.asarray(semipolar_grid, dtype=np.float64)
    internal_intersec, external_intersec, lines = [], [], []
    for l, coords in enumerate(grid):
        line = coords if polyline else coords[[0, -1]]
        list_internal = np.array(polyline_intersections(line, internal_wall)).reshape(-1, 2)
        list_external = np.array(polyline_intersections(line, external_wall)).reshape(-1, 2)
        internal_contact, external_contact = len(list_internal) > 0, len(list_external) > 0
        if not internal_contact and not external_contact:      # :185-186
            continue
        lines.append(l)
        # :199-202 - the EXTERNAL wall's ends are the fallback for both walls
        default_compare_to = np.array([external_wall[0], external_wall[-1]])
        if internal_contact:                                    # :204-212
            compare_to = list_external if external_contact else default_compare_to
            i_min, j_min = argmatrix(distance_matrix(list_internal, compare_to), "min")
            internal_intersec.append(list_internal[i_min])
            if not external_contact:
                external_intersec.append(external_wall[int(-1 * j_min)])
        if external_contact:                                    # :214-221
            compare_to = list_internal if internal_contact else default_compare_to
            i_min, j_min = argmatrix(distance_matrix(list_external, compare_to), "min")
            external_intersec.append(list_external[i_min])
            if not internal_contact:
                internal_intersec.append(internal_wall[int(-1 * j_min)])
    out = (np.array(internal_intersec).reshape(-1, 2), np.array(external_intersec).reshape(-1, 2))
    return out + (np.array(lines, dtype=np.int32),) if return_lines else out


# --------------------------------------------------------------------------- a11: area function
def mid_point(p1, p2):
    """area_function.py:113-121."""
    x1, y1 = p1
    x2, y2 = p2
    return min(x1, x2) + abs(x1 - x2) / 2, min(y1, y2) + abs(y1 - y2) / 2


def area_function(internal_wall, external_wall, alpha=np.pi, beta=2.):
    """area_function.py:124-142 -> (dists (K,), fx (K,)): cumulative distance between the
    mid-points of successive wall pairs, and alpha * (|int - ext| / 2) ** beta."""
    internal_wall = np.asarray(internal_wall, dtype=np.float64)
    external_wall = np.asarray(external_wall, dtype=np.float64)
    assert internal_wall.shape == external_wall.shape
    xs = np.array([mid_point(a, b) for a, b in zip(internal_wall, external_wall)]).reshape(-1, 2)
    radius = np.array([euclidean(a, b) / 2 for a, b in zip(internal_wall, external_wall)])
    fx = alpha * radius ** beta
    d = 0.0
    dists = [d]
    for p1, p2 in zip(xs[:-1], xs[1:]):
        d = euclidean(p1, p2) + d
        dists.append(d)
    return np.array(dists), fx


# --------------------------------------------------------------------------- a12: resampling
def evenly_spaced_fx(x, fx, n_samples=200, dtype=torch.float32):
    """area_function.py:145-159 -> torch.float32 (2, n_samples) (``torch.tensor`` of Python floats;
    ``dtype=torch.float64`` keeps the uncast values for tolerance checks): the abscissae
    linspace(x[0], x[-1], n) and the polyline (x, fx) evaluated there (a vertical line through
    a piecewise-linear graph meets it at the linear interpolant)."""
    x = np.asarray(x, dtype=np.float64)
    fx = np.asarray(fx, dtype=np.float64)
    xs = np.linspace(x[0], x[-1], n_samples)
    out = []
    for xv in xs:
        hit = None
        for s in range(len(x) - 1):
            if x[s] <= xv <= x[s + 1]:
                hit = s
                break
        if hit is None:
            raise ValueError("sample outside the polyline")   # the reference would raise too
        if x[hit + 1] == x[hit] or xv == x[hit]:
            yv = fx[hit]
        elif xv == x[hit + 1]:
            yv = fx[hit + 1]
        else:
            yv = fx[hit] + (xv - x[hit]) / (x[hit + 1] - x[hit]) * (fx[hit + 1] - fx[hit])
        out.append((float(xv), float(yv)))
    return torch.tensor(out, dtype=dtype).T


def vocal_tract_area_function(internal_wall, external_wall, semipolar_grid, alpha=np.pi, beta=2.,
                              n_samples=200):
    """The chain the signatures imply: a10 -> a11 -> a12 for one frame.
    Returns (xfx float64 (2, n_samples) before the float32 cast, K, int_pts, ext_pts, lines)."""
    ii, ee, lines = intersect_semipolar_grid(internal_wall, external_wall, semipolar_grid,
                                             return_lines=True)
    K = len(ii)
    if K < 2:
        return np.full((2, n_samples), np.nan), K, ii, ee, lines
    dists, fx = area_function(ii, ee, alpha, beta)
    xfx = evenly_spaced_fx(dists, fx, n_samples, dtype=torch.float64).numpy()
    return xfx, K, ii, ee, lines
