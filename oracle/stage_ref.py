"""Stage the importable reference files under ``oracle/_ref/`` so that the TIMED CPU baseline on the GPU
box runs the reference's own code instead of a port.

TEST / BENCH INFRASTRUCTURE ONLY.  ``/root/reference`` exists only in the build container; the GPU box
gets a snapshot of this repository.  ``oracle/_ref/`` is git-ignored (reference SOURCES never enter the
history) but not gpurun-ignored, so what this recipe stages travels with the snapshot like the built
``.so`` files do.  The four files below import only torch (plus names that
``oracle.reference_loader`` stubs: six ``vt_tools`` string constants, ``vt_tools.metrics.p2cp_mean``):

    phoneme_to_articulation/encoder_decoder/models.py   ArtSpeech, ArticulatorPredictor
    phoneme_to_articulation/metrics.py                  MeanP2CPDistance, EuclideanDistance
    metrics.py                                          p2cp_distance, euclidean_distance, pearsons_correlation
    tract_variables.py                                  calculate_vocal_tract_variables

Files are copied byte for byte; ``MANIFEST.json`` records their sha256.  Run by
``__graft_entry__.build()`` when the checkout is present: ``python -m oracle.stage_ref``.
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")
FILES = (
    "phoneme_to_articulation/encoder_decoder/models.py",
    "phoneme_to_articulation/metrics.py",
    "metrics.py",
    "tract_variables.py",
)


def stage(reference_root="/root/reference"):
    """Copy FILES from the checkout into oracle/_ref/ (same relative paths).  Returns the directory, or
    None when the checkout is absent (then whatever was staged before stays as it is)."""
    if not os.path.isfile(os.path.join(reference_root, FILES[-1])):
        return REF_DIR if staged() else None
    manifest = {}
    for rel in FILES:
        src, dst = os.path.join(reference_root, rel), os.path.join(REF_DIR, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(src, dst)
        with open(dst, "rb") as f:
            manifest[rel] = hashlib.sha256(f.read()).hexdigest()
    with open(os.path.join(REF_DIR, "MANIFEST.json"), "w") as f:
        json.dump({"source": "vribeiro1/artspeech checkout at " + reference_root, "sha256": manifest}, f, indent=1)
    return REF_DIR


def staged():
    """True if every file is present and matches the manifest."""
    path = os.path.join(REF_DIR, "MANIFEST.json")
    if not os.path.isfile(path):
        return False
    with open(path) as f:
        want = json.load(f)["sha256"]
    for rel in FILES:
        p = os.path.join(REF_DIR, rel)
        if not os.path.isfile(p):
            return False
        with open(p, "rb") as f:
            if hashlib.sha256(f.read()).hexdigest() != want.get(rel):
                return False
    return True


if __name__ == "__main__":
    print(stage())
