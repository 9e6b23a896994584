"""Tract variables with the reference's signature, computed by the CUDA library.

Mirrors tract_variables.py:13-125 of the reference (``ART_SLICES``,
``calculate_vocal_tract_variables``) and adds the batched device entry the evaluation
driver needs: all frames of a batch in one launch instead of one Python call
with three ``.item()`` syncs per search (phoneme_to_articulation/__init__.py:242-261).
"""
from __future__ import annotations

import ctypes

import torch

from . import _lib

# string values of the vt_tools constants imported at tract_variables.py:3-10
LOWER_LIP = "lower-lip"
PHARYNX = "pharynx"
SOFT_PALATE_MIDLINE = "soft-palate-midline"
TONGUE = "tongue"
UPPER_LIP = "upper-lip"
UPPER_INCISOR = "upper-incisor"

# tract_variables.py:13-20
ART_SLICES = {
    "tongue-tip": (30, 45),
    "tongue-body": (10, 30),
    "upper-incisor": (25, 50),
    "hard-palate": (0, 25),
    "soft-palate": (35, 50),
    "velum": (0, 15),
}

# phoneme_to_articulation/__init__.py:37-44
REQUIRED_ARTICULATORS_FOR_TVS = [LOWER_LIP, PHARYNX, SOFT_PALATE_MIDLINE, TONGUE, UPPER_LIP, UPPER_INCISOR]

TV_NAMES = ("LA", "TTCD", "TBCD", "VEL")
# order of art_index_host in the C ABI
_ABI_ORDER = (LOWER_LIP, PHARYNX, SOFT_PALATE_MIDLINE, TONGUE, UPPER_INCISOR, UPPER_LIP)
# (name of arr1, offset of its slice), per TV
_ARR1 = ((LOWER_LIP, 0), (TONGUE, 30), (TONGUE, 10), (SOFT_PALATE_MIDLINE, 0))


def tract_variables_batch(frames, articulators, upper_incisor=None, max_ctas=0):
    """All four TVs for every contour set in one launch.

    frames: (..., N_art, 2, 50) float tensor (CPU or CUDA); ``articulators``: the N_art names in
    row order, or a dict name -> row.  ``upper_incisor``: optional (2, 50) contour shared by all
    sets, or (..., 2, 50) one per set, used instead of a row of ``frames`` (the reference injects
    a fixed reference contour, evaluation.py:93-109).  Returns (val (...,4) fp32, idx (...,4,2) int32,
    poc (...,4,2,2) fp32) on the CUDA device; idx = (i into arr1, j into arr2) of each search,
    poc = (arr1[i], arr2[j]).  ``max_ctas`` > 0 caps the grid (for a launch that runs beside other
    kernels and must leave SMs to them)."""
    if isinstance(articulators, dict):
        index = articulators
    else:
        index = {name: k for k, name in enumerate(articulators)}
    if upper_incisor is not None:
        index = dict(index)
        index.setdefault(UPPER_INCISOR, -1)
    rows = (ctypes.c_int32 * 6)(*[index[name] for name in _ABI_ORDER])  # KeyError like the reference
    if frames.dim() < 3 or frames.shape[-2] != 2 or frames.shape[-1] != 50:
        raise ValueError(f"expected (..., N_art, 2, 50) contours, got {tuple(frames.shape)}")
    dev = _lib.device_of(frames)
    f = frames.detach().to(device=dev, dtype=torch.float32).contiguous()
    lead = f.shape[:-3]
    n_art = f.shape[-3]
    sets = f.numel() // (n_art * 100)
    val = torch.empty(sets, 4, dtype=torch.float32, device=dev)
    idx = torch.empty(sets, 4, 2, dtype=torch.int32, device=dev)
    poc = torch.empty(sets, 4, 2, 2, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        ui, stride = None, 0
        if upper_incisor is not None:
            ui = upper_incisor.detach().to(device=dev, dtype=torch.float32).contiguous()
            if ui.shape[-2:] != (2, 50) or ui.numel() not in (100, sets * 100):
                raise ValueError(f"upper_incisor must be (2, 50) or (..., 2, 50) per set, got {tuple(ui.shape)}")
            stride = 0 if ui.numel() == 100 else 100
        _lib.check(_lib.lib().asb_tract_variables_ex(_lib.ptr(f), sets, n_art, rows,
                                                     _lib.ptr(ui) if ui is not None else None, stride,
                                                     int(max_ctas), _lib.ptr(val), _lib.ptr(idx), _lib.ptr(poc),
                                                     _lib.stream_ptr(dev)))
    return val.view(*lead, 4), idx.view(*lead, 4, 2), poc.view(*lead, 4, 2, 2)


def calculate_vocal_tract_variables(inputs_dict):
    """tract_variables.py:73-125.  ``inputs_dict``: articulator name -> (50, 2) tensor.  Returns
    {"LA"|"TTCD"|"TBCD"|"VEL": {"value": float, "poc_1": (2,) tensor, "poc_2": (2,) tensor},
     "LP"|"TTCL"|"TBCL"|"GLO": None}; the PoCs are views into the inputs like the reference's."""
    arrs = [inputs_dict[name] for name in _ABI_ORDER]  # KeyError for a missing articulator
    for a in arrs:
        if a.dim() != 2 or a.shape != (50, 2):
            raise ValueError(f"expected (50, 2) contours, got {tuple(a.shape)}")
    frame = torch.stack([a.transpose(0, 1) for a in arrs], dim=0)  # (6, 2, 50)
    val, idx, _ = tract_variables_batch(frame, {name: k for k, name in enumerate(_ABI_ORDER)})
    val = val.cpu().tolist()
    idx = idx.cpu().tolist()

    uinc, spm = inputs_dict[UPPER_INCISOR], inputs_dict[SOFT_PALATE_MIDLINE]

    def arr2_point(k, j):
        if k == 0:
            return inputs_dict[UPPER_LIP][j]
        if k == 1:
            return uinc[25 + j]
        if k == 2:
            return uinc[j] if j < 25 else spm[35 + (j - 25)]
        return inputs_dict[PHARYNX][j]

    out = {"LA": None, "LP": None, "TTCD": None, "TTCL": None, "TBCD": None, "TBCL": None, "VEL": None,
           "GLO": None}
    for k, name in enumerate(TV_NAMES):
        a1_name, a1_off = _ARR1[k]
        i, j = idx[k]
        out[name] = {"value": val[k], "poc_1": inputs_dict[a1_name][a1_off + i], "poc_2": arr2_point(k, j)}
    return out


def tract_variables(sentences_ids, frame_ids, outputs, targets, lengths, phonemes, articulators, save_to):
    """The frame driver of the reference, phoneme_to_articulation/__init__.py:201-297, with the same
    arguments and the same ``{save_to}/{sentence_id}/tract_variables.csv`` files (columns: sentence,
    frame, phoneme, then value + two PoCs per TV for the target, then for the prediction).

    The reference calls ``calculate_vocal_tract_variables`` twice per frame from Python (about
    0.35 ms each, three ``.item()`` syncs per search); here the tract variables of every frame of
    the batch, prediction and target, are two kernel launches and one device->host copy.
    outputs/targets: (bs, seq_len, n_articulators, 2, 50), CPU or CUDA."""
    import os

    import pandas as pd
    articulators = list(articulators)
    val_p, _, poc_p = tract_variables_batch(outputs, articulators)
    val_t, _, poc_t = tract_variables_batch(targets, articulators)
    val_p, poc_p, val_t, poc_t = (t.cpu().tolist() for t in (val_p, poc_p, val_t, poc_t))
    for b, (sentence_id, length, sentence_frames, sentence_phonemes) in enumerate(
            zip(sentences_ids, lengths, frame_ids, phonemes)):
        saves_sentence_dir = os.path.join(save_to, sentence_id)
        os.makedirs(saves_sentence_dir, exist_ok=True)
        rows = []
        for t, (frame, phoneme) in enumerate(zip(sentence_frames, sentence_phonemes)):
            if t >= int(length):
                break
            item = {"sentence": sentence_id, "frame": frame, "phoneme": phoneme}
            for which, val, poc in (("target", val_t, poc_t), ("pred", val_p, poc_p)):   # :263-285
                for k, name in enumerate(TV_NAMES):
                    item[f"{name}_{which}"] = val[b][t][k]
                    item[f"{name}_{which}_poc_1_x"] = poc[b][t][k][0][0]
                    item[f"{name}_{which}_poc_1_y"] = poc[b][t][k][0][1]
                    item[f"{name}_{which}_poc_2_x"] = poc[b][t][k][1][0]
                    item[f"{name}_{which}_poc_2_y"] = poc[b][t][k][1][1]
            rows.append(item)
        pd.DataFrame(rows).to_csv(os.path.join(saves_sentence_dir, "tract_variables.csv"), index=False)
