"""Oracle restatement of the contour geometry: P2CP, MED, Pearson, tract variables.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

Per SURVEY.md F2 the parity truth is the reference algorithm evaluated on
float64 inputs (``dtype=torch.float64``, the default here): the reference's
own fp32 ``torch.cdist`` takes the matmul formulation for > 25 points and is
noisier (1.6e-3 max relative) than the 1e-5 gate.  Distances here use the
direct sqrt(dx^2 + dy^2) form, which is what cdist computes up to rounding.
"""
from __future__ import annotations

import torch

# tract_variables.py:13-20
ART_SLICES = {
    "tongue-tip": (30, 45),
    "tongue-body": (10, 30),
    "upper-incisor": (25, 50),
    "hard-palate": (0, 25),
    "soft-palate": (35, 50),
    "velum": (0, 15),
}
# string values of the vt_tools constants imported at tract_variables.py:3-10
LOWER_LIP = "lower-lip"
PHARYNX = "pharynx"
SOFT_PALATE_MIDLINE = "soft-palate-midline"
TONGUE = "tongue"
UPPER_LIP = "upper-lip"
UPPER_INCISOR = "upper-incisor"
REQUIRED = [LOWER_LIP, PHARYNX, SOFT_PALATE_MIDLINE, TONGUE, UPPER_LIP, UPPER_INCISOR]
TV_NAMES = ("LA", "TTCD", "TBCD", "VEL")


def _pairwise(a, b):
    """(*, n, 2), (*, m, 2) -> (*, n, m) Euclidean distances (what torch.cdist returns)."""
    d = a.unsqueeze(-2) - b.unsqueeze(-3)
    return torch.sqrt((d * d).sum(-1))


def mean_p2cp(u, v, reduction="none", dtype=torch.float64):
    """MeanP2CPDistance.forward, phoneme_to_articulation/metrics.py:32-46."""
    u = u.to(dtype); v = v.to(dtype)
    n, m = u.shape[-2], v.shape[-2]
    dist = _pairwise(u, v)                                   # :41
    u2cp = dist.min(dim=-1).values                           # :42
    v2cp = dist.min(dim=-2).values                           # :43
    res = (u2cp.sum(-1) / n + v2cp.sum(-1) / m) / 2          # :44
    fn = getattr(torch, reduction, lambda t: t)              # :30
    return fn(res)


def p2cp_distance(outputs, targets, dtype=torch.float64):
    """metrics.py:38-52: (B,T,N,2,D) x2 -> (B,T,N)."""
    return mean_p2cp(outputs.transpose(-1, -2), targets.transpose(-1, -2), "none", dtype)


def euclidean_distance(outputs, targets, dtype=torch.float64):
    """metrics.py:54-68 over phoneme_to_articulation/metrics.py:10-24: (B,T,N,2,D) -> (B,T,N)."""
    o = outputs.to(dtype); t = targets.to(dtype)
    d = torch.sqrt((o[..., 0, :] - t[..., 0, :]) ** 2 + (o[..., 1, :] - t[..., 1, :]) ** 2)
    return d.mean(dim=-1)


def pearsons_correlation(outputs, targets, dtype=torch.float64):
    """metrics.py:9-35, INCLUDING its quirk: the x target mean is taken from the
    x *outputs* (:22).  (B,T,N,2,D) -> x_corr, y_corr each (B,N,D)."""
    eps = 1e-5
    o = outputs.to(dtype); t = targets.to(dtype)
    xo, yo = o[:, :, :, 0, :], o[:, :, :, 1, :]
    xt, yt = t[:, :, :, 0, :], t[:, :, :, 1, :]
    vxo = xo - xo.mean(dim=1, keepdim=True)
    vxt = xt - xo.mean(dim=1, keepdim=True)      # quirk kept (:22)
    x_corr = (vxo * vxt).sum(1) / (torch.sqrt((vxo ** 2).sum(1)) * torch.sqrt((vxt ** 2).sum(1)) + eps)
    vyo = yo - yo.mean(dim=1, keepdim=True)
    vyt = yt - yt.mean(dim=1, keepdim=True)
    y_corr = (vyo * vyt).sum(1) / (torch.sqrt((vyo ** 2).sum(1)) * torch.sqrt((vyt ** 2).sum(1)) + eps)
    return x_corr, y_corr


def closest_pair(arr1, arr2):
    """_calculate_TV, tract_variables.py:23-35.  arr1 (P,2), arr2 (R,2).
    Returns (value, i, j): min over dim0 first, then over the column minima, so
    ties resolve to the smallest j and, for that j, the smallest i."""
    dist = _pairwise(arr1, arr2)                 # :24
    min_d0, argmin_d0 = dist.min(dim=0)          # :26
    min_d1, argmin_d1 = min_d0.min(dim=0)        # :27
    j = int(argmin_d1)
    i = int(argmin_d0[j])
    return float(min_d1), i, j


def tv_operands(frame, art_index):
    """Point sets of the four TVs for one frame (N_art, 2, D) -> list of (arr1, arr2),
    each (P,2); tract_variables.py:38-70.  ``art_index`` maps the six required
    names to rows of ``frame``; rows are (2, D) and are transposed to (D, 2) as the
    driver does (phoneme_to_articulation/__init__.py:253-260)."""
    g = lambda name: frame[art_index[name]].transpose(0, 1)
    llip, ulip = g(LOWER_LIP), g(UPPER_LIP)
    tongue, uinc = g(TONGUE), g(UPPER_INCISOR)
    spm, phar = g(SOFT_PALATE_MIDLINE), g(PHARYNX)
    s = ART_SLICES
    palate = torch.cat([uinc[slice(*s["hard-palate"])], spm[slice(*s["soft-palate"])]], dim=0)
    return [
        (llip, ulip),                                                            # LA   :38-41
        (tongue[slice(*s["tongue-tip"])], uinc[slice(*s["upper-incisor"])]),     # TTCD :44-51
        (tongue[slice(*s["tongue-body"])], palate),                              # TBCD :54-62
        (spm[slice(*s["velum"])], phar),                                         # VEL  :65-70
    ]


def tract_variables_batch(frames, art_index, dtype=torch.float64):
    """calculate_vocal_tract_variables (tract_variables.py:73-125) over F frames.
    frames: (F, N_art, 2, D).  Returns values (F,4) float64, idx (F,4,2) int32
    [i into arr1, j into arr2], poc (F,4,2,2) [poc_1 xy, poc_2 xy] in ``dtype``."""
    frames = frames.to(dtype)
    Fn = frames.shape[0]
    val = torch.zeros(Fn, 4, dtype=torch.float64)
    idx = torch.zeros(Fn, 4, 2, dtype=torch.int32)
    poc = torch.zeros(Fn, 4, 2, 2, dtype=dtype)
    for f in range(Fn):
        for k, (a1, a2) in enumerate(tv_operands(frames[f], art_index)):
            v, i, j = closest_pair(a1, a2)
            val[f, k] = v
            idx[f, k, 0] = i; idx[f, k, 1] = j
            poc[f, k, 0] = a1[i]; poc[f, k, 1] = a2[j]
    return val, idx, poc


def tract_variables_batch_vectorized(frames, art_index, dtype=torch.float64):
    """Same result as ``tract_variables_batch`` computed with batched torch ops (used to
    check larger cases quickly; equality with the loop version is itself tested)."""
    frames = frames.to(dtype)
    fr = frames.transpose(-1, -2)  # (F, N, D, 2)
    g = lambda name: fr[:, art_index[name]]
    s = ART_SLICES
    llip, ulip, tongue = g(LOWER_LIP), g(UPPER_LIP), g(TONGUE)
    uinc, spm, phar = g(UPPER_INCISOR), g(SOFT_PALATE_MIDLINE), g(PHARYNX)
    palate = torch.cat([uinc[:, slice(*s["hard-palate"])], spm[:, slice(*s["soft-palate"])]], dim=1)
    ops = [(llip, ulip),
           (tongue[:, slice(*s["tongue-tip"])], uinc[:, slice(*s["upper-incisor"])]),
           (tongue[:, slice(*s["tongue-body"])], palate),
           (spm[:, slice(*s["velum"])], phar)]
    Fn = frames.shape[0]
    val = torch.zeros(Fn, 4, dtype=torch.float64)
    idx = torch.zeros(Fn, 4, 2, dtype=torch.int32)
    poc = torch.zeros(Fn, 4, 2, 2, dtype=dtype)
    ar = torch.arange(Fn)
    for k, (a1, a2) in enumerate(ops):
        dist = _pairwise(a1, a2)                       # (F, P, R)
        min_d0, arg_d0 = dist.min(dim=1)               # over arr1 -> (F, R)
        min_d1, arg_d1 = min_d0.min(dim=1)             # over arr2 -> (F,)
        j = arg_d1
        i = arg_d0[ar, j]
        val[:, k] = min_d1
        idx[:, k, 0] = i.to(torch.int32); idx[:, k, 1] = j.to(torch.int32)
        poc[:, k, 0] = a1[ar, i]; poc[:, k, 1] = a2[ar, j]
    return val, idx, poc


def calculate_vocal_tract_variables(inputs_dict, dtype=torch.float64):
    """Dict-in/dict-out form, tract_variables.py:73-125.  Values are (D,2) tensors."""
    names = sorted(inputs_dict.keys())
    art_index = {n: k for k, n in enumerate(names)}
    frame = torch.stack([inputs_dict[n].to(dtype).transpose(0, 1) for n in names], dim=0)
    val, idx, poc = tract_variables_batch(frame[None], art_index, dtype)
    out = {"LP": None, "TTCL": None, "TBCL": None, "GLO": None}
    for k, name in enumerate(TV_NAMES):
        out[name] = {"value": float(val[0, k]), "poc_1": poc[0, k, 0], "poc_2": poc[0, k, 1]}
    return out


CRITICAL_TV_TO_ARTICULATORS = {
    "LA": ("lower-lip", "upper-lip"), "TTCD": ("tongue", "upper-incisor"),
    "TBCD": ("tongue", "upper-incisor"), "VEL": ("soft-palate", "pharynx"),
}


def critical_loss(output_shapes, reference_arrays, critical_mask, TVs, articulators, dtype=torch.float64):
    """CriticalLoss.forward, phoneme_to_articulation/principal_components/losses.py:52-99 (no
    denormalisation): mean over the entries with critical_mask == 1 of the minimum distance
    between the two (whole) articulators of each tract variable.  Differentiable torch ops, so the
    tests can take its gradient as the truth for the CUDA path's backward."""
    TVs = sorted(TVs)                                                       # :38
    articulators = list(articulators)
    x = output_shapes.to(dtype)
    if "upper-incisor" not in articulators:                                 # :39-42, :64-70
        articulators = sorted(articulators + ["upper-incisor"])
        k = articulators.index("upper-incisor")
        x = torch.cat([x[:, :, :k], reference_arrays.to(dtype), x[:, :, k:]], dim=2)
    index = {a: i for i, a in enumerate(articulators)}
    per_tv = []
    for tv in TVs:                                                          # :74-93
        a1, a2 = CRITICAL_TV_TO_ARTICULATORS[tv]
        d = _pairwise(x[..., index[a1], :, :].transpose(-1, -2), x[..., index[a2], :, :].transpose(-1, -2))
        per_tv.append(d.flatten(-2).min(dim=-1).values)                     # (B, T)  :95-98
    loss = torch.stack(per_tv, dim=1)                                       # (B, Ntvs, T)
    return loss[critical_mask == 1].mean()                                  # :99
