"""Seeded synthetic inputs of the shapes the evaluation path consumes.

There is no dataset on the box, so bench.py, smoke() and the tests all draw
tokens / contours from here.  Everything is generated on the CPU with an explicit
``torch.Generator`` so that the same seed gives the same bytes on every machine.
"""
from __future__ import annotations

import math

import torch

# Sorted articulator names of the thesis configs
# (thesis_config/model_free_phoneme_to_articulation/test_model_free.yaml:8-18) plus the
# upper incisor that evaluation.py:93-109 injects at its sorted position.
ARTICULATORS = [
    "arytenoid-cartilage", "epiglottis", "lower-incisor", "lower-lip", "pharynx",
    "soft-palate-midline", "thyroid-cartilage", "tongue", "upper-lip", "vocal-folds",
]
TV_ARTICULATORS = sorted(ARTICULATORS + ["upper-incisor"])
TV_ART_INDEX = {name: TV_ARTICULATORS.index(name) for name in
                ("lower-lip", "pharynx", "soft-palate-midline", "tongue", "upper-incisor", "upper-lip")}
UPPER_INCISOR_SLOT = TV_ARTICULATORS.index("upper-incisor")

# The fixed upper-incisor contour the synthesis script hard-codes
# (generate_vocal_tract_shape_v2.py:42-67, SynthesisDataset.ref_array); the only real
# contour data in the reference repository.
UPPER_INCISOR_REF = torch.tensor([
    [0.4690135, 0.4662627, 0.46351194, 0.46076113, 0.45801032, 0.45527017, 0.45255125,
     0.44987226, 0.44734472, 0.44498724, 0.44288498, 0.4411229, 0.43968427, 0.43846822,
     0.43745804, 0.43643433, 0.43517762, 0.43364525, 0.43158084, 0.42894173, 0.42584795,
     0.42241973, 0.4186992, 0.4149394, 0.41118246, 0.4073549, 0.4033833, 0.3991887,
     0.39429778, 0.38863158, 0.38197833, 0.37412614, 0.36517772, 0.35574925, 0.3459434,
     0.33627468, 0.3272577, 0.31895584, 0.3117499, 0.30570334, 0.30080932, 0.29706097,
     0.2944497, 0.29292405, 0.29247546, 0.29290503, 0.29401374, 0.29574504, 0.29775932, 0.3],
    [0.2479401, 0.24794108, 0.24794206, 0.24794301, 0.247944, 0.24794495, 0.24794582,
     0.24794653, 0.24794671, 0.24794629, 0.24794495, 0.24794239, 0.24793896, 0.2479367,
     0.24793586, 0.2479395, 0.24795061, 0.24796885, 0.24799258, 0.24802142, 0.24802947,
     0.24799073, 0.24789768, 0.24770513, 0.24740562, 0.24718466, 0.24722782, 0.2476354,
     0.24900925, 0.25144967, 0.2552429, 0.26067525, 0.26764584, 0.27554935, 0.28428498,
     0.29333425, 0.30217865, 0.31071305, 0.3183063, 0.32485318, 0.33003068, 0.33351567,
     0.33525217, 0.3349038, 0.33241457, 0.3281533, 0.3224889, 0.3155383, 0.308003, 0.3],
], dtype=torch.float32)


def tokens(batch, seq_len, vocab_size=45, seed=0):
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, vocab_size, (batch, seq_len), generator=g, dtype=torch.int64)


def sorted_lengths(batch, lo, hi, seed=0):
    """Descending lengths in [lo, hi] with the first equal to ``hi`` (so max(lengths) == hi)."""
    g = torch.Generator().manual_seed(seed + 17)
    L = torch.randint(lo, hi + 1, (batch,), generator=g, dtype=torch.int64)
    L[0] = hi
    return torch.sort(L, descending=True).values.to(torch.int32)


def contours(n_frames, n_articulators=10, n_samples=50, seed=0, noise=0.002, device=None):
    """(F, N, 2, D) fp32 arcs: centre U(0.3,0.7)^2, radius U(0.05,0.15), phase U(0,2pi),
    point k at angle phase + 3k/(D-1), plus N(0, noise^2).  ``device``: generate there with that
    device's generator (the million-frame workloads of bench.py; same law, other bytes)."""
    dev = torch.device("cpu") if device is None else torch.device(device)
    g = torch.Generator(device=dev).manual_seed(seed)
    F, N, D = n_frames, n_articulators, n_samples
    c = 0.3 + 0.4 * torch.rand(F, N, 2, 1, generator=g, device=dev)
    r = 0.05 + 0.10 * torch.rand(F, N, 1, 1, generator=g, device=dev)
    ph = 2 * math.pi * torch.rand(F, N, 1, 1, generator=g, device=dev)
    k = torch.arange(D, dtype=torch.float32, device=dev).view(1, 1, 1, D)
    ang = ph + 3.0 * k / (D - 1)
    pts = torch.cat([torch.cos(ang), torch.sin(ang)], dim=2)
    out = c + r * pts + noise * torch.randn(F, N, 2, D, generator=g, device=dev)
    return out.to(torch.float32).contiguous()


def perturbed(target, seed=1, sigma=0.01):
    """A stand-in 'prediction' = target + N(0, sigma^2) when no model is run (on the target's device)."""
    g = torch.Generator(device=target.device).manual_seed(seed)
    noise = torch.randn(target.shape, generator=g, device=target.device)
    return (target + sigma * noise).to(torch.float32).contiguous()


def inject_upper_incisor(x):
    """(…, 10, 2, D) -> (…, 11, 2, D) with the fixed upper-incisor contour inserted at its
    sorted slot, as evaluation.py:93-109 does with ``reference_arrays``."""
    ref = UPPER_INCISOR_REF.to(x.device, x.dtype).expand(*x.shape[:-3], 1, 2, UPPER_INCISOR_REF.shape[-1])
    s = UPPER_INCISOR_SLOT
    return torch.cat([x[..., :s, :, :], ref, x[..., s:, :, :]], dim=-3).contiguous()


def tv_frames(n_frames, seed=0):
    """(F, 11, 2, 50) contour sets with the upper incisor in place."""
    return inject_upper_incisor(contours(n_frames, 10, 50, seed))


# ---- area-function inputs ---------------------------------------------------
# Default grid of the synthetic area-function workload: build_semipolar_grid((0.5, 0.5), 0, 0,
# 0.01, pi / 36) -> 117 lines (50 larynx + 17 polar + 50 mouth, area_function.py:83-108).
AREA_GRID_ARGS = dict(center=(0.5, 0.5), theta_rad=0.0, omega_rad=0.0, linear_step=0.01,
                      polar_step_rad=math.pi / 36)
_MOUTH_LEN, _ARC_LEN, _LARYNX_LEN = 0.47, 0.22, 0.47


def _tract_path(u, rho, center=(0.5, 0.5)):
    """Point at path coordinate u and distance rho from the grid's internal side: mouth section
    (lines hanging from y = cy), quarter fan around the centre, larynx section (lines leaving
    x = cx); follows the geometry of area_function.py:44-79 for theta = omega = 0."""
    cx, cy = center
    a = (_MOUTH_LEN - u).clamp(min=0.0)
    phi = ((u - _MOUTH_LEN) / _ARC_LEN).clamp(0.0, 1.0) * (math.pi / 2)
    b = (u - _MOUTH_LEN - _ARC_LEN).clamp(min=0.0)
    x = cx - a + rho * torch.sin(phi)
    y = cy - rho * torch.cos(phi) + b
    return x, y


def walls(n_frames, n_points=50, seed=0, noise=0.0005, device=None):
    """Internal and external vocal-tract walls, two (F, n_points, 2) fp32 tensors: a smooth open
    curve through the grid region at depth rho(u) = U(0.10,0.16) + U(0.01,0.04) sin(...), the
    external wall offset outward by U(0.02,0.05) +- 0.01 (so 0.01..0.06), both ends of each wall
    trimmed independently by U(0, 0.08) so that some grid lines meet only one wall, plus
    N(0, noise^2)."""
    dev = torch.device("cpu") if device is None else torch.device(device)
    g = torch.Generator(device=dev).manual_seed(seed)
    F = n_frames
    total = _MOUTH_LEN + _ARC_LEN + _LARYNX_LEN
    r = lambda lo, hi: lo + (hi - lo) * torch.rand(F, 1, generator=g, dtype=torch.float64, device=dev)
    rho0, amp, freq, ph = r(0.10, 0.16), r(0.01, 0.04), r(0.5, 2.0), r(0.0, 2 * math.pi)
    off0, oamp, ofreq, oph = r(0.02, 0.05), r(0.0, 0.01), r(0.5, 3.0), r(0.0, 2 * math.pi)
    k = torch.arange(n_points, dtype=torch.float64, device=dev).view(1, -1) / (n_points - 1)
    out = []
    for is_ext in (False, True):
        u0, u1 = r(0.0, 0.08), total - r(0.0, 0.08)
        u = u0 + (u1 - u0) * k
        rho = rho0 + amp * torch.sin(2 * math.pi * freq * u / total + ph)
        if is_ext:
            rho = rho + off0 + oamp * torch.sin(2 * math.pi * ofreq * u / total + oph)
        x, y = _tract_path(u, rho)
        pts = torch.stack([x, y], dim=-1)
        pts = pts + noise * torch.randn(pts.shape, generator=g, dtype=torch.float64, device=dev)
        out.append(pts.to(torch.float32).contiguous())
    return out[0], out[1]


# ---- a stand-in for the reference's DataLoader (dataset.py collate_fn tuples) ----------------------
class _Config:
    RES, PIXEL_SPACING = 136, 1.412          # what dataset_config carries for run_test's to_mm factor


class _Dataset:
    dataset_config = _Config()


class SyntheticLoader(list):
    """A list of the 8-tuples the reference's collate_fn yields - (sentences_ids, sentences, targets,
    lengths, phonemes, reference_arrays, sentence_frames, voicing) - with a ``.dataset.dataset_config``,
    which is all ``run_test`` (evaluation.py:17-160) asks of its dataloader."""
    dataset = _Dataset()


def loader(n_batches=2, batch=3, seq_len=7, vocab=45, seed=0):
    out = SyntheticLoader()
    for k in range(n_batches):
        L = sorted_lengths(batch, 2, seq_len, seed=seed + k)
        tok = tokens(batch, seq_len, vocab, seed=seed + 10 * k)
        tgt = contours(batch * seq_len, 10, 50, seed=seed + 100 + k).view(batch, seq_len, 10, 2, 50)
        ids = [f"S{k}_{b}" for b in range(batch)]
        frames = [["%04d" % (50 * b + t) for t in range(int(L[b]))] for b in range(batch)]
        phon = [[chr(97 + (t % 26)) for t in range(int(L[b]))] for b in range(batch)]
        ref = UPPER_INCISOR_REF.view(1, 1, 1, 2, 50).expand(batch, seq_len, 1, 2, 50).contiguous()
        out.append((ids, tok, tgt, L, phon, ref, frames, None))
    return out
