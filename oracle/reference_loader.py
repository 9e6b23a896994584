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


def available():
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "tract_variables.py"))


def _load(name, relpath):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REFERENCE_ROOT, relpath))
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


def models():
    """reference phoneme_to_articulation/encoder_decoder/models.py"""
    return _load("_ref_models", "phoneme_to_articulation/encoder_decoder/models.py")


def p2a_metrics():
    """reference phoneme_to_articulation/metrics.py (MeanP2CPDistance, EuclideanDistance)"""
    return _load("_ref_p2a_metrics", "phoneme_to_articulation/metrics.py")


def tract_variables():
    """reference tract_variables.py with the vt_tools constant stub"""
    _stub_vt_tools()
    return _load("_ref_tract_variables", "tract_variables.py")
