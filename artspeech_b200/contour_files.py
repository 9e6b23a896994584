"""The reference's on-disk contour format and the per-frame error report over it.

``save_outputs`` (phoneme_to_articulation/__init__.py:143-199) writes, per sentence, one ``.npy``
per frame and articulator - ``contours/{frame}_{art}.npy`` (prediction) and
``contours/{frame}_{art}_true.npy`` (target), each a (2, 50) array - and
``report_phoneme_to_articulation.py:198-239`` reads them back frame by frame, builds (1, 1, N, 2, 50)
tensors and calls ``p2cp_distance`` / ``euclidean_distance`` once per frame.  Here a sentence is read
into one (F, N, 2, 50) pair of tensors and both metrics are ONE kernel launch each over all its
frames (``asb_p2cp``, ``asb_mean_euclidean``); the rows produced are the reference's
``metrics_data`` items, field for field.  B-spline regularisation (``regularize_Bsplines``, :186) is
the un-vendored ``vt_tools`` package and is not restated: arrays are written as given.
"""
from __future__ import annotations

import os

import numpy as np
import torch

from .metrics import euclidean_distance, p2cp_distance


def frame_name(frame):
    """Frame ids are written as given by ``save_outputs`` (strings from the dataset) and looked up
    as ``"%04d" % int(frame)`` by the report (report_phoneme_to_articulation.py:201)."""
    return frame if isinstance(frame, str) else "%04d" % int(frame)


def save_contours(saves_sentence_dir, frames, outputs, targets, articulators):
    """outputs/targets (F, N, 2, D) tensors or arrays for the F frames of one sentence ->
    ``{saves_sentence_dir}/contours/{frame}_{art}[_true].npy`` (__init__.py:181-195; the i-th array
    belongs to ``sorted(articulators)[i]``)."""
    cdir = os.path.join(saves_sentence_dir, "contours")
    os.makedirs(cdir, exist_ok=True)
    out = outputs.detach().cpu().numpy() if torch.is_tensor(outputs) else np.asarray(outputs)
    tgt = targets.detach().cpu().numpy() if torch.is_tensor(targets) else np.asarray(targets)
    for f, frame in enumerate(frames):
        for i_art, art in enumerate(sorted(articulators)):
            with open(os.path.join(cdir, f"{frame_name(frame)}_{art}.npy"), "wb") as fh:
                np.save(fh, out[f, i_art])
            with open(os.path.join(cdir, f"{frame_name(frame)}_{art}_true.npy"), "wb") as fh:
                np.save(fh, tgt[f, i_art])


def load_sentence_contours(contours_dir, frames, articulators):
    """All frames of one sentence -> (pred, true) float32 tensors (F, N, 2, D) in pinned host memory
    (articulators in the given order, report_phoneme_to_articulation.py:205-213)."""
    first = np.load(os.path.join(contours_dir, f"{frame_name(frames[0])}_{articulators[0]}.npy"))
    shape = (len(frames), len(articulators)) + tuple(first.shape)
    pin = torch.cuda.is_available()
    pred = torch.empty(shape, dtype=torch.float32, pin_memory=pin)
    true = torch.empty(shape, dtype=torch.float32, pin_memory=pin)
    pn, tn = pred.numpy(), true.numpy()
    for f, frame in enumerate(frames):
        stem = os.path.join(contours_dir, frame_name(frame))
        for a, art in enumerate(articulators):
            pn[f, a] = np.load(f"{stem}_{art}.npy")          # .type(torch.float), :218-219
            tn[f, a] = np.load(f"{stem}_{art}_true.npy")
    return pred, true


def sentence_error_rows(sentence_name, contours_dir, frames, phonemes, articulators, to_mm):
    """The ``metrics_data`` items of one sentence (report_phoneme_to_articulation.py:198-237):
    one row per (frame, articulator) with p2cp, p2cp_mm, euclidean, euclidean_mm."""
    pred, true = load_sentence_contours(contours_dir, frames, articulators)
    dev = torch.device("cuda", torch.cuda.current_device())
    pd_, td_ = pred.to(dev, non_blocking=True).unsqueeze(0), true.to(dev, non_blocking=True).unsqueeze(0)
    p2cp = p2cp_distance(pd_, td_)[0].cpu().tolist()                 # (F, N)
    med = euclidean_distance(pd_, td_)[0].cpu().tolist()
    rows = []
    for f, (frame, phoneme) in enumerate(zip(frames, phonemes)):
        for i, articulator in enumerate(articulators):
            rows.append({
                "sentence_name": sentence_name, "frame": frame, "phoneme": phoneme, "articulator": articulator,
                "p2cp": p2cp[f][i], "p2cp_mm": p2cp[f][i] * to_mm,
                "euclidean": med[f][i], "euclidean_mm": med[f][i] * to_mm,
            })
    return rows
