"""The committed golden recipe (``python -m oracle.make_golden``) runs END TO END in one process and
reproduces every committed fixture bit for bit (build container only: it executes the unmodified
reference from /root/reference)."""
import filecmp
import json
import os

import numpy as np
import pytest

from oracle import reference_loader as ref

pytestmark = pytest.mark.skipif(not ref.available(), reason="reference checkout not present")


def test_all_loaders_share_one_process():
    """Every loader in either order: the vt_tools.metrics stub carries what each of them imports."""
    a = ref.root_metrics()
    b = ref.area_function()
    c = ref.root_metrics()
    assert hasattr(a, "pearsons_correlation") and hasattr(c, "p2cp_distance")
    assert abs(b.euclidean((0.0, 0.0), (3.0, 4.0)) - 5.0) < 1e-15


def test_make_golden_reproduces_committed_fixtures(tmp_path, golden_dir):
    from oracle import make_golden
    make_golden.main(str(tmp_path))
    produced = sorted(os.listdir(tmp_path))
    committed = sorted(f for f in os.listdir(golden_dir) if f in produced)
    assert produced == committed, (produced, committed)
    assert len(produced) >= 14
    for name in produced:
        got, want = os.path.join(tmp_path, name), os.path.join(golden_dir, name)
        if name.endswith(".npz"):
            g, w = np.load(got), np.load(want)
            assert sorted(g.files) == sorted(w.files), name
            for k in w.files:
                assert g[k].dtype == w[k].dtype and g[k].shape == w[k].shape, (name, k)
                assert np.array_equal(g[k], w[k], equal_nan=g[k].dtype.kind == "f"), (name, k)
        elif name.endswith(".json"):
            with open(got) as f1, open(want) as f2:
                assert json.load(f1) == json.load(f2), name
        else:
            assert filecmp.cmp(got, want, shallow=False), name
