"""Import the UNMODIFIED reference modules by file path (build container only).

TEST INFRASTRUCTURE ONLY.  ``/root/reference`` does not exist on the GPU box, so
nothing that runs there may call this; it is used by ``oracle/make_golden.py`` and
by CPU tests that are skipped when the checkout is absent.

Mechanics: ``models.py`` and ``phoneme_to_articulation/metrics.py``
import only torch and load as they are; ``tract_variables.py`` needs six name
constants from the un-vendored ``vt_tools`` package, supplied by an in-process stub
(they are only dict keys, so any consistent strings give identical numerics).  The
``phoneme_to_articulation`` package itself is never imported (its __init__ needs
vt_tools / vt_shape_gen / pandas).
"""
from __future__ import annotations

import importlib.util
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("ARTSPEECH_REFERENCE", "/root/reference")
STAGED_ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


def available():
    """The full checkout is present (build container)."""
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "tract_variables.py"))


def baseline_available():
    """The files the timed CPU baseline needs are present: the checkout, or the copies staged under
    oracle/_ref/ by oracle/stage_ref.py (those travel to the GPU box)."""
    if available():
        return True
    from oracle import stage_ref
    return stage_ref.staged()


def _root():
    return REFERENCE_ROOT if available() else STAGED_ROOT


def _load(name, relpath):
    spec = importlib.util.spec_from_file_location(name, os.path.join(_root(), relpath))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _stub_vt_tools():
    if "vt_tools" in sys.modules:
        return
    vt = types.ModuleType("vt_tools")
    vt.LOWER_LIP = "lower-lip"
    vt.PHARYNX = "pharynx"
    vt.SOFT_PALATE_MIDLINE = "soft-palate-midline"
    vt.TONGUE = "tongue"
    vt.UPPER_LIP = "upper-lip"
    vt.UPPER_INCISOR = "upper-incisor"
    sys.modules["vt_tools"] = vt


def _stub_vt_metrics():
    """ONE ``vt_tools.metrics`` stub shared by every loader (the package is un-vendored): each loader
    needs different names from it, so all of them are always present - ``p2cp_mean`` (imported and
    unused by root metrics.py:4), ``euclidean`` = sqrt(dx^2 + dy^2) (its published definition; used by
    area_function.py:128,139) and ``distance_matrix`` (used only next to GEOS calls, which cannot run
    here).  Loaders may be called in any order in one process."""
    import numpy as np
    _stub_vt_tools()
    vt = sys.modules["vt_tools"]
    vm = sys.modules.get("vt_tools.metrics")
    if vm is None:
        vm = types.ModuleType("vt_tools.metrics")
        sys.modules["vt_tools.metrics"] = vm
    if getattr(vm, "euclidean", None) is None:
        vm.euclidean = lambda u, v: np.sqrt((u[0] - v[0]) ** 2 + (u[1] - v[1]) ** 2)
    for name in ("p2cp_mean", "distance_matrix"):
        if not hasattr(vm, name):
            setattr(vm, name, None)
    vt.metrics = vm
    return vm


def models():
    """reference phoneme_to_articulation/encoder_decoder/models.py"""
    return _load("_ref_models", "phoneme_to_articulation/encoder_decoder/models.py")


def p2a_metrics():
    """reference phoneme_to_articulation/metrics.py (MeanP2CPDistance, EuclideanDistance)"""
    return _load("_ref_p2a_metrics", "phoneme_to_articulation/metrics.py")


def root_metrics():
    """reference metrics.py (pearsons_correlation, p2cp_distance, euclidean_distance).  Its imports of
    the un-vendored ``vt_tools.metrics.p2cp_mean`` (unused by these functions) and of the
    ``phoneme_to_articulation`` package (whose __init__ needs vt_tools / vt_shape_gen) are stubbed:
    the package stub only carries the real phoneme_to_articulation/metrics.py module."""
    _stub_vt_metrics()
    if "phoneme_to_articulation" not in sys.modules:
        pkg = types.ModuleType("phoneme_to_articulation")
        pkg.__path__ = []
        sys.modules["phoneme_to_articulation"] = pkg
        sys.modules["phoneme_to_articulation.metrics"] = p2a_metrics()
    return _load("_ref_root_metrics", "metrics.py")


def tract_variables():
    """reference tract_variables.py with the vt_tools constant stub"""
    _stub_vt_tools()
    return _load("_ref_tract_variables", "tract_variables.py")


def area_function():
    """reference area_function.py with its absent imports stubbed:
    ``funcy.lmap`` = list(map(...)); ``shapely.geometry`` = empty placeholder classes (so only the
    functions that never touch GEOS can run: rotate, build_semipolar_grid, mid_point, argmatrix,
    area_function); ``vt_tools.metrics.euclidean`` = sqrt(dx^2 + dy^2) (its published definition,
    the package itself is un-vendored); ``np.float`` aliased to ``float`` for :130 (NumPy >= 1.24)."""
    import numpy as np
    _stub_vt_metrics()
    if "funcy" not in sys.modules:
        fy = types.ModuleType("funcy")
        fy.lmap = lambda fn, seq: list(map(fn, seq))
        sys.modules["funcy"] = fy
    if "shapely" not in sys.modules:
        sh = types.ModuleType("shapely")
        sg = types.ModuleType("shapely.geometry")
        sg.LineString = type("LineString", (), {})
        sg.Point = type("Point", (), {})
        sh.geometry = sg
        sys.modules["shapely"] = sh
        sys.modules["shapely.geometry"] = sg
    if not hasattr(np, "float"):
        np.float = float
    return _load("_ref_area_function", "area_function.py")


def critical_loss():
    """reference ``CriticalLoss`` (phoneme_to_articulation/principal_components/losses.py:23-99).
    The module's other imports (helpers, the principal_components models, vt_tools) cannot be
    satisfied here, so ONLY the class statement is compiled - its source text unmodified, cut out
    of the file with ``ast`` - in a namespace holding torch, nn and the six vt_tools name
    constants it uses as dict keys (``SOFT_PALATE`` = "soft-palate")."""
    import ast
    import torch
    import torch.nn as nn
    path = os.path.join(REFERENCE_ROOT, "phoneme_to_articulation/principal_components/losses.py")
    with open(path) as f:
        src = f.read()
    tree = ast.parse(src)
    node = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "CriticalLoss")
    code = compile(ast.Module(body=[node], type_ignores=[]), path, "exec")
    ns = {"torch": torch, "nn": nn, "LOWER_LIP": "lower-lip", "PHARYNX": "pharynx", "SOFT_PALATE": "soft-palate",
          "TONGUE": "tongue", "UPPER_INCISOR": "upper-incisor", "UPPER_LIP": "upper-lip"}
    exec(code, ns)
    return ns["CriticalLoss"]


def tract_variables_driver():
    """reference frame driver ``tract_variables(...)`` (phoneme_to_articulation/__init__.py:201-297).
    The package's __init__ cannot be imported (vt_tools / vt_shape_gen), so ONLY this function's
    statement is compiled - source text unmodified, cut out with ``ast`` - in a namespace holding
    ``os``, ``pd`` and the reference's own ``calculate_vocal_tract_variables``."""
    import ast
    import pandas as pd
    path = os.path.join(REFERENCE_ROOT, "phoneme_to_articulation/__init__.py")
    with open(path) as f:
        tree = ast.parse(f.read())
    node = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "tract_variables")
    code = compile(ast.Module(body=[node], type_ignores=[]), path, "exec")
    ns = {"os": os, "pd": pd, "calculate_vocal_tract_variables": tract_variables().calculate_vocal_tract_variables}
    exec(code, ns)
    return ns["tract_variables"]


def _cut(relpath, name, ns, kind="function"):
    """Compile one top-level statement (function or class ``name``) of a reference file, source text
    unmodified, in namespace ``ns`` - for modules whose other imports cannot be satisfied here."""
    import ast
    path = os.path.join(REFERENCE_ROOT, relpath)
    with open(path) as f:
        tree = ast.parse(f.read())
    want = ast.FunctionDef if kind == "function" else ast.ClassDef
    node = next(n for n in tree.body if isinstance(n, want) and n.name == name)
    exec(compile(ast.Module(body=[node], type_ignores=[]), path, "exec"), ns)
    return ns[name]


def run_test():
    """reference ``run_test`` (phoneme_to_articulation/encoder_decoder/evaluation.py:17-160) assembled
    from the reference's own pieces: its metrics.py functions, ``make_padding_mask`` (helpers.py:79-91),
    ``tract_variables`` and ``save_outputs`` (phoneme_to_articulation/__init__.py:128-297) - each cut out
    as an unmodified statement because the enclosing modules import un-vendored packages."""
    import numpy as np
    import pandas as pd
    import torch
    from tqdm import tqdm
    rm = root_metrics()
    tv = tract_variables()
    ns_pkg = {"os": os, "pd": pd, "np": np,
              "calculate_vocal_tract_variables": tv.calculate_vocal_tract_variables}
    drv = _cut("phoneme_to_articulation/__init__.py", "tract_variables", ns_pkg)
    save = _cut("phoneme_to_articulation/__init__.py", "save_outputs", ns_pkg)
    mask = _cut("helpers.py", "make_padding_mask", {"torch": torch})
    ns = {"np": np, "os": os, "torch": torch, "tqdm": tqdm, "make_padding_mask": mask,
          "pearsons_correlation": rm.pearsons_correlation, "p2cp_distance": rm.p2cp_distance,
          "euclidean_distance": rm.euclidean_distance, "save_outputs": save, "tract_variables": drv,
          "REQUIRED_ARTICULATORS_FOR_TVS": ["lower-lip", "pharynx", "soft-palate-midline", "tongue", "upper-lip",
                                            "upper-incisor"],
          "UPPER_INCISOR": "upper-incisor"}
    return _cut("phoneme_to_articulation/encoder_decoder/evaluation.py", "run_test", ns)


def principal_components_models():
    """reference ``PrincipalComponentsPredictor`` / ``PrincipalComponentsArtSpeech``
    (phoneme_to_articulation/principal_components/models/rnn.py:11-109): the two class statements,
    unmodified, with ``RNNType`` (phoneme_to_articulation/__init__.py:47-49) and
    ``make_indices_dict`` (helpers.py:94-114) cut out of their own files the same way."""
    from enum import Enum
    from functools import reduce

    import torch
    import torch.nn as nn
    from torch.nn.utils.rnn import pack_padded_sequence, pad_packed_sequence
    rnn_type = _cut("phoneme_to_articulation/__init__.py", "RNNType", {"Enum": Enum, "nn": nn}, kind="class")
    mid = _cut("helpers.py", "make_indices_dict", {})
    ns = {"torch": torch, "nn": nn, "reduce": reduce, "pack_padded_sequence": pack_padded_sequence,
          "pad_packed_sequence": pad_packed_sequence, "make_indices_dict": mid, "RNNType": rnn_type}
    rel = "phoneme_to_articulation/principal_components/models/rnn.py"
    _cut(rel, "PrincipalComponentsPredictor", ns, kind="class")
    return _cut(rel, "PrincipalComponentsArtSpeech", ns, kind="class"), rnn_type
