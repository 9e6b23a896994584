"""On-disk contour format (save_outputs, phoneme_to_articulation/__init__.py:181-195) and the error
report rows over it (report_phoneme_to_articulation.py:198-237)."""
import os

import numpy as np
import pytest
import torch

from artspeech_b200 import contour_files, synthetic

ARTS = ["arytenoid-cartilage", "lower-lip", "tongue"]


def _sentence(tmp_path):
    F = 9
    tgt = synthetic.contours(F, 3, 50, seed=41)
    pred = tgt + 0.01 * torch.randn(tgt.shape, generator=torch.Generator().manual_seed(1))
    frames = ["%04d" % (100 + 3 * k) for k in range(F)]
    contour_files.save_contours(str(tmp_path), frames, pred, tgt, list(reversed(ARTS)))
    return frames, pred, tgt


def test_file_names_and_round_trip(tmp_path):
    frames, pred, tgt = _sentence(tmp_path)
    cdir = os.path.join(str(tmp_path), "contours")
    names = sorted(os.listdir(cdir))
    assert len(names) == 2 * 9 * 3 and "0100_lower-lip.npy" in names and "0124_tongue_true.npy" in names
    one = np.load(os.path.join(cdir, "0103_tongue.npy"))
    assert one.shape == (2, 50) and one.dtype == np.float32
    assert np.array_equal(one, pred[1, 2].numpy())               # i-th array <-> sorted(articulators)[i]
    # the report addresses frames as "%04d" % int(frame)
    p, t = contour_files.load_sentence_contours(cdir, [int(f) for f in frames], ARTS)
    assert torch.equal(p, pred) and torch.equal(t, tgt)
    p2, _ = contour_files.load_sentence_contours(cdir, frames, ["tongue", "lower-lip"])
    assert torch.equal(p2[:, 0], pred[:, 2]) and torch.equal(p2[:, 1], pred[:, 1])


@pytest.mark.gpu
def test_error_rows_match_per_frame_oracle(tmp_path):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from oracle import geometry as og
    frames, pred, tgt = _sentence(tmp_path)
    phonemes = ["a"] * len(frames)
    rows = contour_files.sentence_error_rows("s1", os.path.join(str(tmp_path), "contours"), frames, phonemes,
                                             ARTS, to_mm=1.5)
    assert len(rows) == len(frames) * len(ARTS)
    assert list(rows[0].keys()) == ["sentence_name", "frame", "phoneme", "articulator", "p2cp", "p2cp_mm",
                                    "euclidean", "euclidean_mm"]
    k = 0
    for f, frame in enumerate(frames):                            # the reference's per-frame loop
        p = og.p2cp_distance(pred[f][None, None], tgt[f][None, None])[0, 0]
        e = og.euclidean_distance(pred[f][None, None], tgt[f][None, None])[0, 0]
        for i, art in enumerate(ARTS):
            r = rows[k]; k += 1
            assert (r["sentence_name"], r["frame"], r["articulator"]) == ("s1", frame, art)
            assert abs(r["p2cp"] - p[i].item()) <= 1e-5 * p[i].item()
            assert abs(r["euclidean"] - e[i].item()) <= 1e-5 * e[i].item()
            assert r["p2cp_mm"] == r["p2cp"] * 1.5 and r["euclidean_mm"] == r["euclidean"] * 1.5
