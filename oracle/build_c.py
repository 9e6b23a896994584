"""Compile the plain-C oracle (oracle/area_c.c) with gcc and bind it with ctypes.

TEST INFRASTRUCTURE ONLY.  ``build()`` is called by ``__graft_entry__.build()`` ("building the
checker is not using it"); the resulting oracle/_build/libarea_oracle.so is git-ignored and
travels to the GPU box with the snapshot.  ``-ffp-contract=off`` keeps every product and sum
separately rounded, which is what the NumPy restatement and the CUDA kernel's fp64 path do.
"""
from __future__ import annotations

import ctypes
import os
import shutil
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "area_c.c")
OUT_DIR = os.path.join(HERE, "_build")
LIB = os.path.join(OUT_DIR, "libarea_oracle.so")
_handle = None


def build(force=False):
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= os.path.getmtime(SRC):
        return LIB
    gcc = shutil.which("gcc")
    if gcc is None:
        if os.path.exists(LIB):
            return LIB
        raise RuntimeError("gcc not found and oracle/_build/libarea_oracle.so is not prebuilt")
    os.makedirs(OUT_DIR, exist_ok=True)
    subprocess.check_call([gcc, "-O2", "-ffp-contract=off", "-fopenmp", "-fPIC", "-shared", "-std=c11",
                           SRC, "-o", LIB, "-lm"])
    return LIB


def _lib():
    global _handle
    if _handle is None:
        h = ctypes.CDLL(build())
        h.area_oracle_batch.restype = ctypes.c_int
        h.area_oracle_batch.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int,
                                        ctypes.c_void_p, ctypes.c_int, ctypes.c_double, ctypes.c_double,
                                        ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                        ctypes.c_void_p, ctypes.c_int]
        _handle = h
    return _handle


def area_function_batch(walls_int, walls_ext, grid_ends, alpha=np.pi, beta=2.0, n_samples=200, threads=1,
                        want_points=True):
    """walls_* (F, Nw, 2) float32 arrays, grid_ends (n_lines, 2, 2) float64.
    Returns dict(xfx (F,2,n_samples) f64, K (F) i32, pts (F,n_lines,2,2) f64, lines (F,n_lines) i32)."""
    wi = np.ascontiguousarray(np.asarray(walls_int), dtype=np.float32)
    we = np.ascontiguousarray(np.asarray(walls_ext), dtype=np.float32)
    ends = np.ascontiguousarray(grid_ends, dtype=np.float64)
    assert wi.shape == we.shape and wi.ndim == 3 and wi.shape[2] == 2
    F, nw, _ = wi.shape
    n_lines = ends.shape[0]
    xfx = np.empty((F, 2, n_samples), dtype=np.float64)
    K = np.empty((F,), dtype=np.int32)
    pts = np.empty((F, n_lines, 2, 2), dtype=np.float64) if want_points else None
    lines = np.empty((F, n_lines), dtype=np.int32) if want_points else None
    rc = _lib().area_oracle_batch(wi.ctypes.data, we.ctypes.data, F, nw, ends.ctypes.data, n_lines,
                                  float(alpha), float(beta), int(n_samples), xfx.ctypes.data, K.ctypes.data,
                                  pts.ctypes.data if want_points else None,
                                  lines.ctypes.data if want_points else None, int(threads))
    if rc != 0:
        raise ValueError("area_oracle_batch: bad arguments")
    return {"xfx": xfx, "K": K, "pts": pts, "lines": lines}


if __name__ == "__main__":
    print(build(force=True))
