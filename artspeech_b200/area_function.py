"""Area function with the reference's signatures, computed by the CUDA library.

Mirrors area_function.py of the reference: ``rotate`` :12-28, ``build_semipolar_grid`` :31-110,
``mid_point`` :113-121, ``area_function`` :124-142, ``evenly_spaced_fx`` :145-159, ``argmatrix``
:162-172, ``intersect_semipolar_grid`` :175-223 - NumPy arrays in, NumPy arrays / a torch tensor out -
and adds the batched device entry ``area_function_batch``: walls of all frames ->
resampled area profiles in one launch instead of 2 * n_lines + 200 shapely/GEOS calls per frame.

The grid is configuration, built once on the host exactly as the reference builds it and uploaded
as line end points (row a9 of SURVEY.md 8a).  Everything per frame runs in the CUDA library; there
is no CPU fallback (``_lib`` raises if the library or a CUDA device is missing).
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib


# ------------------------------------------------------------------ host-side configuration (a9)
def rotate(point, ang_rad):
    """area_function.py:12-28: [[cos, sin], [-sin, cos]] @ point."""
    c, s = np.cos(ang_rad), np.sin(ang_rad)
    return np.matmul(np.array([[c, s], [-s, c]]), point)


def build_semipolar_grid(center, theta_rad, omega_rad, linear_step, polar_step_rad, grid_res=50):
    """Maeda's semipolar grid, area_function.py:31-110 -> (n_lines, grid_res, 2) float64:
    larynx lines (last first), polar lines (last first), mouth lines; every line is
    ``grid_res`` samples from its internal to its external end."""
    center = np.asarray(center, dtype=np.float64)
    ends = []  # (internal end, external end) per line, in output order
    larynx = [(rotate(np.array([0., y]), omega_rad) + center, rotate(np.array([0.4, y]), omega_rad) + center)
              for y in np.arange(0., 0.5, linear_step)]
    angles = np.arange(theta_rad - polar_step_rad, -(np.pi / 2) + omega_rad, -polar_step_rad)
    polar = [(np.zeros(2) + center, rotate(np.array([0., -0.4]), a) + center) for a in angles]
    mouth = [(rotate(np.array([x, 0.]), theta_rad) + center, rotate(np.array([x, -0.4]), theta_rad) + center)
             for x in np.arange(0., -0.5, -linear_step)]
    ends.extend(reversed(larynx))
    ends.extend(reversed(polar))
    ends.extend(mouth)
    grid = np.empty((len(ends), grid_res, 2), dtype=np.float64)
    for l, (p_int, p_ext) in enumerate(ends):
        grid[l, :, 0] = np.linspace(p_int[0], p_ext[0], grid_res)
        grid[l, :, 1] = np.linspace(p_int[1], p_ext[1], grid_res)
    return grid


def mid_point(p1, p2):
    """area_function.py:113-121 (host helper; the kernels compute it inline)."""
    (x1, y1), (x2, y2) = p1, p2
    return min(x1, x2) + abs(x1 - x2) / 2, min(y1, y2) + abs(y1 - y2) / 2


def argmatrix(mtx, minmax):
    """area_function.py:162-172: (row, column) of the first min/max in row-major order."""
    assert minmax in ["min", "max"]
    flat = getattr(np, f"arg{minmax}")(mtx)
    _, m = mtx.shape
    return flat // m, flat % m


def grid_endpoints(semipolar_grid):
    """(n_lines, grid_res, 2) grid -> (n_lines, 2, 2) float64 [line][internal|external][x|y]."""
    g = np.asarray(semipolar_grid, dtype=np.float64)
    if g.ndim != 3 or g.shape[-1] != 2 or g.shape[1] < 2:
        raise ValueError(f"expected a (n_lines, grid_res, 2) grid, got {g.shape}")
    return np.ascontiguousarray(g[:, [0, -1], :])


class DeviceGrid:
    """Grid end points resident on a CUDA device (upload once, reuse for every batch)."""

    def __init__(self, semipolar_grid, device=None):
        ends = grid_endpoints(semipolar_grid)
        dev = device if device is not None else _lib.device_of()
        self.n_lines = ends.shape[0]
        self.ends = torch.from_numpy(ends).to(dev)


def _as_device_grid(grid, dev):
    if isinstance(grid, DeviceGrid):
        return grid
    return DeviceGrid(grid, dev)


# ------------------------------------------------------------------ batched device entry
def area_function_batch(walls_int, walls_ext, grid, alpha=np.pi, beta=2., n_samples=200, return_points=False):
    """Frames of wall pairs -> resampled area functions, one launch.

    walls_int / walls_ext: (..., Nw, 2) float tensors (CPU or CUDA; computed from their fp32
    values); grid: a ``DeviceGrid`` or a (n_lines, grid_res, 2) array.  Returns a dict of CUDA tensors:
    ``xfx`` (..., 2, n_samples) fp32 (NaN rows where fewer than two grid lines touch a wall), ``K`` (...)
    int32 touched grid lines, and with ``return_points`` ``int_pts`` / ``ext_pts`` (..., n_lines, 2) fp64
    (first K rows valid) and ``lines`` (..., n_lines) int32 (-1 padded)."""
    if walls_int.shape != walls_ext.shape or walls_int.dim() < 2 or walls_int.shape[-1] != 2:
        raise ValueError(f"expected two (..., Nw, 2) walls, got {tuple(walls_int.shape)} and {tuple(walls_ext.shape)}")
    dev = _lib.device_of(walls_int, walls_ext)
    wi = walls_int.detach().to(device=dev, dtype=torch.float32).contiguous()
    we = walls_ext.detach().to(device=dev, dtype=torch.float32).contiguous()
    lead, nw = wi.shape[:-2], wi.shape[-2]
    frames = wi.numel() // (2 * nw)
    g = _as_device_grid(grid, dev)
    xfx = torch.empty(frames, 2, n_samples, dtype=torch.float32, device=dev)
    K = torch.empty(frames, dtype=torch.int32, device=dev)
    pts = lines = None
    if return_points:
        pts = torch.empty(frames, g.n_lines, 2, 2, dtype=torch.float64, device=dev)
        lines = torch.empty(frames, g.n_lines, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        _lib.check(_lib.lib().asb_area_function(
            _lib.ptr(wi), _lib.ptr(we), frames, nw, _lib.ptr(g.ends), g.n_lines, float(alpha), float(beta),
            int(n_samples), _lib.ptr(xfx), _lib.ptr(K), _lib.ptr(pts) if return_points else None,
            _lib.ptr(lines) if return_points else None, _lib.stream_ptr(dev)))
    out = {"xfx": xfx.view(*lead, 2, n_samples), "K": K.view(lead)}
    if return_points:
        out["int_pts"] = pts[:, :, 0, :].reshape(*lead, g.n_lines, 2)
        out["ext_pts"] = pts[:, :, 1, :].reshape(*lead, g.n_lines, 2)
        out["lines"] = lines.view(*lead, g.n_lines)
    return out


# ------------------------------------------------------------------ the reference's per-frame API
def intersect_semipolar_grid(internal_wall, external_wall, semipolar_grid):
    """area_function.py:175-223: walls (Nw, 2) and the grid -> (internal_intersec (K, 2),
    external_intersec (K, 2)) float64 arrays, one pair per grid line that touches a wall."""
    wi = torch.as_tensor(np.asarray(internal_wall))
    we = torch.as_tensor(np.asarray(external_wall))
    r = area_function_batch(wi, we, semipolar_grid, n_samples=2, return_points=True)
    K = int(r["K"].item())
    return r["int_pts"][:K].cpu().numpy(), r["ext_pts"][:K].cpu().numpy()


def area_function(internal_wall, external_wall, alpha=np.pi, beta=2.):
    """area_function.py:124-142: paired wall points (K, 2) x 2 -> (dists (K,), fx (K,)) float64."""
    internal_wall, external_wall = np.asarray(internal_wall), np.asarray(external_wall)
    assert internal_wall.shape == external_wall.shape
    if internal_wall.ndim != 2 or internal_wall.shape[1] != 2 or internal_wall.shape[0] < 1:
        raise ValueError(f"expected (K, 2) wall points, got {internal_wall.shape}")
    dev = _lib.device_of()
    K = internal_wall.shape[0]
    pi = torch.from_numpy(np.ascontiguousarray(internal_wall, dtype=np.float64)).to(dev)
    pe = torch.from_numpy(np.ascontiguousarray(external_wall, dtype=np.float64)).to(dev)
    dists = torch.empty(K, dtype=torch.float64, device=dev)
    fx = torch.empty(K, dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        _lib.check(_lib.lib().asb_area_from_pairs(_lib.ptr(pi), _lib.ptr(pe), 1, K, float(alpha), float(beta),
                                                  _lib.ptr(dists), _lib.ptr(fx), _lib.stream_ptr(dev)))
    return dists.cpu().numpy(), fx.cpu().numpy()


def evenly_spaced_fx(x, fx, n_samples=200):
    """area_function.py:145-159: -> torch.float32 (2, n_samples) CPU tensor [abscissae; fx there]."""
    x = np.ascontiguousarray(x, dtype=np.float64)
    fx = np.ascontiguousarray(fx, dtype=np.float64)
    if x.ndim != 1 or x.shape != fx.shape:
        raise ValueError(f"expected two (K,) arrays, got {x.shape} and {fx.shape}")
    if x.shape[0] < 2:
        raise ValueError("evenly_spaced_fx needs at least two points")  # shapely raises on a 1-point LineString
    dev = _lib.device_of()
    xd, fd = torch.from_numpy(x).to(dev), torch.from_numpy(fx).to(dev)
    out = torch.empty(2, n_samples, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        _lib.check(_lib.lib().asb_evenly_spaced_fx(_lib.ptr(xd), _lib.ptr(fd), 1, x.shape[0], int(n_samples),
                                                   _lib.ptr(out), _lib.stream_ptr(dev)))
    return out.cpu()
