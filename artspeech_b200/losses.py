"""``CriticalLoss`` of the reference (phoneme_to_articulation/principal_components/losses.py:23-99)
on the closest-pair kernel of the CUDA library (``asb_min_distance``).

The reference materialises, per tract variable, the full (B, T, D, D) ``torch.cdist`` matrix and
takes its minimum; here one warp per (frame, tract variable) searches the D x D pairs in registers
and returns the minimum and the pair that attains it.  The loss is differentiable: the gradient
of a minimum distance flows to the two points of the closest pair only, ``d|p - q|/dp =
(p - q)/|p - q|`` (zero when the points coincide, as in cdist's backward), which is what autograd
computes for the reference's cdist -> min graph.  There is no CPU implementation.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from . import _lib

LOWER_LIP, PHARYNX, SOFT_PALATE, TONGUE, UPPER_INCISOR, UPPER_LIP = (
    "lower-lip", "pharynx", "soft-palate", "tongue", "upper-incisor", "upper-lip")


def min_distance(u, v):
    """u (*, 2, n), v (*, 2, m) float32 CUDA tensors -> (dist (*), ij (*, 2) int32)."""
    if u.shape[:-2] != v.shape[:-2] or u.shape[-2] != 2 or v.shape[-2] != 2:
        raise ValueError(f"expected (*, 2, n) and (*, 2, m) with equal leading dims, got {tuple(u.shape)}, "
                         f"{tuple(v.shape)}")
    dev = _lib.device_of(u, v)
    uc = u.detach().to(device=dev, dtype=torch.float32).contiguous()
    vc = v.detach().to(device=dev, dtype=torch.float32).contiguous()
    lead = u.shape[:-2]
    items = uc.numel() // (2 * u.shape[-1]) if uc.numel() else 0
    d = torch.empty(items, dtype=torch.float32, device=dev)
    ij = torch.empty(items, 2, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        _lib.check(_lib.lib().asb_min_distance(_lib.ptr(uc), _lib.ptr(vc), items, u.shape[-1], v.shape[-1],
                                               _lib.ptr(d), _lib.ptr(ij), _lib.stream_ptr(dev)))
    return d.view(lead), ij.view(*lead, 2)


class _MinDistance(torch.autograd.Function):
    @staticmethod
    def forward(ctx, u, v):
        d, ij = min_distance(u, v)
        ctx.save_for_backward(u, v, d, ij)
        return d

    @staticmethod
    def backward(ctx, grad):
        u, v, d, ij = ctx.saved_tensors
        i = ij[..., 0].long().unsqueeze(-1).unsqueeze(-1).expand(*ij.shape[:-1], 2, 1)
        j = ij[..., 1].long().unsqueeze(-1).unsqueeze(-1).expand(*ij.shape[:-1], 2, 1)
        p = torch.gather(u, -1, i)                      # (*, 2, 1) the closest point of u
        q = torch.gather(v, -1, j)
        dd = d.unsqueeze(-1).unsqueeze(-1)
        g = torch.where(dd > 0, (p - q) / dd, torch.zeros_like(p)) * grad.unsqueeze(-1).unsqueeze(-1)
        gu = torch.zeros_like(u).scatter_(-1, i, g)
        gv = torch.zeros_like(v).scatter_(-1, j, -g)
        return gu, gv


class CriticalLoss(nn.Module):
    """losses.py:23-99: mean over the critical (frame, TV) entries of the minimum distance between
    the two articulators of each tract variable.  Same constructor and ``forward`` arguments."""

    TV_TO_ARTICULATOR_MAP = {
        "LA": [LOWER_LIP, UPPER_LIP],
        "TTCD": [TONGUE, UPPER_INCISOR],
        "TBCD": [TONGUE, UPPER_INCISOR],
        "VEL": [SOFT_PALATE, PHARYNX],
    }

    def __init__(self, TVs, articulators, denormalize_fn=None):
        super().__init__()
        self.TVs = sorted(TVs)
        self.inject_reference = UPPER_INCISOR not in articulators
        if UPPER_INCISOR not in articulators:
            articulators = sorted(articulators + [UPPER_INCISOR])
        self.articulators_indices = {articulator: i for i, articulator in enumerate(articulators)}
        self.denorm_fn = denormalize_fn

    def forward(self, output_shapes, target_shapes, reference_arrays, critical_mask):
        if len(self.TVs) == 0:
            return torch.tensor(0, device=target_shapes.device, dtype=torch.float)
        if self.inject_reference:                                          # losses.py:64-70
            ref_index = self.articulators_indices[UPPER_INCISOR]
            output_shapes = torch.concat([output_shapes[:, :, :ref_index], reference_arrays,
                                          output_shapes[:, :, ref_index:]], dim=2)
        dists = []
        for TV in self.TVs:                                                # losses.py:74-93
            arrays = []
            for articulator in self.TV_TO_ARTICULATOR_MAP[TV]:
                a = output_shapes[..., self.articulators_indices[articulator], :, :]
                if self.denorm_fn and articulator != UPPER_INCISOR:
                    a = self.denorm_fn[articulator](a)
                arrays.append(a)
            dists.append(_MinDistance.apply(arrays[0], arrays[1]))         # (B, T)
        critical_loss = torch.stack(dists, dim=1)                          # (B, Ntvs, T), losses.py:95-98
        return critical_loss[critical_mask == 1].mean()
