"""Build libartspeech_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python -m artspeech_b200.build [--force]
"""
from __future__ import annotations

import glob
import hashlib
import os
import shutil
import subprocess
import sys

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
LIB_PATH = os.path.join(PKG_DIR, "libartspeech_b200.so")
OBJ_DIR = os.path.join(PKG_DIR, "csrc", "build")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
]


def _nvcc():
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found; cannot build libartspeech_b200.so")
    return exe


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def headers():
    return sorted(glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(PKG_DIR, "..", "include", "*.h")))


def _digest(paths, extra=()):
    h = hashlib.sha256()
    for item in extra:
        h.update(str(item).encode())
    for path in paths:
        h.update(os.path.basename(path).encode())
        with open(path, "rb") as f:
            h.update(f.read())
    return h.hexdigest()[:16]


def source_hash():
    """Hash of everything the library is compiled from (sources, headers, flags).  It is compiled INTO
    the library (asb_build_hash) and compared at load time, so a stale prebuilt .so is never used."""
    return _digest(sources() + headers(), NVCC_FLAGS)


def build_library(force=False, verbose=False):
    """Compile every csrc/*.cu whose content hash (source + headers + flags) differs from the one
    recorded beside its object, and link the shared library.  Staleness is decided by CONTENT, not by
    mtimes: objects and the .so travel to the GPU box with the snapshot and must match the sources."""
    os.makedirs(OBJ_DIR, exist_ok=True)
    hdrs = headers()
    nvcc = _nvcc()
    whole = source_hash()
    objs = []
    procs = []
    for src in sources():
        name = os.path.basename(src)[:-3]
        obj = os.path.join(OBJ_DIR, name + ".o")
        tag_path = obj + ".hash"
        # asb_error.cu carries the hash of the WHOLE source set, so it is rebuilt whenever anything changes
        flags = list(NVCC_FLAGS) + ([f'-DASB_SOURCE_HASH="{whole}"'] if name == "asb_error" else [])
        tag = _digest([src] + hdrs, flags)
        objs.append(obj)
        have = open(tag_path).read().strip() if os.path.exists(tag_path) and os.path.exists(obj) else None
        if force or have != tag:
            cmd = [nvcc, *flags, "-Xptxas", "-v", "-c", src, "-o", obj]
            procs.append((src, tag_path, tag,
                          subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for src, tag_path, tag, p in procs:
        out, _ = p.communicate()
        if verbose or p.returncode != 0:
            sys.stderr.write(out)
        if p.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}")
        with open(tag_path, "w") as f:
            f.write(tag + "\n")
    if force or procs or not os.path.exists(LIB_PATH) or library_hash() != whole:
        cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB_PATH, *objs, "-lcudart"]
        subprocess.check_call(cmd)
    if library_hash() != whole:
        raise RuntimeError(f"{LIB_PATH} does not carry the hash of the current sources ({whole})")
    return LIB_PATH


def library_hash():
    """asb_build_hash() of the library on disk (None if it is missing or too old to have one).  Read in a
    child process, because a dlopen'ed library cannot be replaced in this one."""
    if not os.path.exists(LIB_PATH):
        return None
    code = ("import ctypes,sys; l=ctypes.CDLL(sys.argv[1]); f=l.asb_build_hash; f.restype=ctypes.c_char_p; "
            "print(f().decode())")
    r = subprocess.run([sys.executable, "-c", code, LIB_PATH], capture_output=True, text=True)
    return r.stdout.strip() if r.returncode == 0 else None


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose=True))
