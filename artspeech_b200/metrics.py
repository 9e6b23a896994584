"""Contour metrics with the reference's names and signatures, computed by the CUDA library.

Mirrors (reference paths):
  * ``MeanP2CPDistance``, ``EuclideanDistance``   phoneme_to_articulation/metrics.py:5-46
  * ``p2cp_distance``, ``euclidean_distance``, ``pearsons_correlation``   metrics.py:9-68
  * ``P2CPDistance``                              phoneme_to_articulation/encoder_decoder/metrics.py:7-26

CPU tensors are accepted like in the reference (its evaluation loop passes CPU tensors,
evaluation.py:65-72): they are copied to the GPU, the kernel runs there and the result comes back
on the input's device.  There is no CPU implementation.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from . import _lib


def _as_f32_cuda(t, dev):
    return t.detach().to(device=dev, dtype=torch.float32, non_blocking=True)


def _p2cp_raw(u, v, items, n, m, layout, dev):
    out = torch.empty(items, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        _lib.check(_lib.lib().asb_p2cp(_lib.ptr(u), _lib.ptr(v), _lib.ptr(out), items, n, m, layout,
                                       _lib.stream_ptr(dev)))
    return out


def _is_planar_view(t):
    """True if ``t`` (*, n, 2) is a transposed view of contiguous (*, 2, n) storage, i.e. what
    ``p2cp_distance`` passes down (metrics.py:47-50)."""
    return t.dim() >= 2 and t.shape[-1] == 2 and t.transpose(-1, -2).is_contiguous()


class MeanP2CPDistance(nn.Module):
    """phoneme_to_articulation/metrics.py:27-46.  u_ (*, N, 2), v_ (*, M, 2) -> (*) or reduced."""

    def __init__(self, reduction="mean"):
        super().__init__()
        self.reduction = getattr(torch, reduction, lambda x: x)

    def forward(self, u_, v_):
        if u_.shape[-1] != 2 or v_.shape[-1] != 2 or u_.shape[:-2] != v_.shape[:-2]:
            raise ValueError(f"expected (*, N, 2) and (*, M, 2) with equal leading dims, got "
                             f"{tuple(u_.shape)} and {tuple(v_.shape)}")
        n, m = u_.shape[-2], v_.shape[-2]
        lead = u_.shape[:-2]
        dev = _lib.device_of(u_, v_)
        if _is_planar_view(u_) and _is_planar_view(v_):
            layout = _lib.LAYOUT_PLANAR
            uc = _as_f32_cuda(u_.transpose(-1, -2), dev)
            vc = _as_f32_cuda(v_.transpose(-1, -2), dev)
        else:
            layout = _lib.LAYOUT_INTERLEAVED
            uc = _as_f32_cuda(u_, dev).contiguous()
            vc = _as_f32_cuda(v_, dev).contiguous()
        items = 1
        for s in lead:
            items *= int(s)
        out = _p2cp_raw(uc, vc, items, n, m, layout, dev).view(lead)
        out = out.to(device=u_.device, dtype=u_.dtype if u_.dtype.is_floating_point else torch.float32)
        return self.reduction(out)


class _PointDistance(torch.autograd.Function):
    """Per-point distance of planar contours (asb_point_distance) with its gradient wrt the first argument
    (asb_point_distance_backward); the targets of the criterion carry no gradient."""

    @staticmethod
    def forward(ctx, o, t):
        n = o.shape[-1]
        items = o.numel() // (2 * n)
        out = torch.empty(*o.shape[:-2], n, dtype=torch.float32, device=o.device)
        with torch.cuda.device(o.device):
            _lib.check(_lib.lib().asb_point_distance(_lib.ptr(o), _lib.ptr(t), _lib.ptr(out), items, n,
                                                     _lib.stream_ptr(o.device)))
        ctx.save_for_backward(o, t)
        return out

    @staticmethod
    def backward(ctx, grad_out):
        o, t = ctx.saved_tensors
        n = o.shape[-1]
        items = o.numel() // (2 * n)
        g = grad_out.to(torch.float32).contiguous()
        grad_o = torch.empty_like(o)
        with torch.cuda.device(o.device):
            _lib.check(_lib.lib().asb_point_distance_backward(_lib.ptr(o), _lib.ptr(t), _lib.ptr(g), _lib.ptr(grad_o),
                                                              items, n, _lib.stream_ptr(o.device)))
        return grad_o, None


class EuclideanDistance(nn.Module):
    """phoneme_to_articulation/metrics.py:5-24: per-point distance (*, N_art, N_samples) of (*, N_art, 2,
    N_samples) contours; the training criterion (train_phoneme_to_articulation.py:86), differentiable wrt the
    outputs.  Evaluated by the library on the CUDA device (CPU inputs are moved there and the result back)."""

    def __init__(self, reduction="mean"):
        super().__init__()
        self.reduction = getattr(torch, reduction, lambda x: x)

    def forward(self, outputs, targets):
        if outputs.shape != targets.shape or outputs.shape[-2] != 2:
            raise ValueError("expected two (*, 2, N_samples) tensors of equal shape")
        dev = _lib.device_of(outputs, targets)
        o = outputs.to(device=dev, dtype=torch.float32).contiguous()
        t = targets.detach().to(device=dev, dtype=torch.float32).contiguous()
        return self.reduction(_PointDistance.apply(o, t)).to(outputs.device)


def p2cp_distance(outputs, targets):
    """metrics.py:38-52.  (bs, seq_len, N_art, 2, N_samples) x2 -> (bs, seq_len, N_art)."""
    return MeanP2CPDistance(reduction="none")(outputs.transpose(-1, -2), targets.transpose(-1, -2))


def euclidean_distance(outputs, targets):
    """metrics.py:54-68.  (bs, seq_len, N_art, 2, N_samples) x2 -> (bs, seq_len, N_art)."""
    if outputs.shape != targets.shape or outputs.shape[-2] != 2:
        raise ValueError("expected two (*, 2, N_samples) tensors of equal shape")
    dev = _lib.device_of(outputs, targets)
    n = outputs.shape[-1]
    lead = outputs.shape[:-2]
    o = _as_f32_cuda(outputs, dev).contiguous()
    t = _as_f32_cuda(targets, dev).contiguous()
    items = o.numel() // (2 * n)
    out = torch.empty(items, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        _lib.check(_lib.lib().asb_mean_euclidean(_lib.ptr(o), _lib.ptr(t), _lib.ptr(out), items, n,
                                                 _lib.stream_ptr(dev)))
    return out.view(lead).to(device=outputs.device, dtype=outputs.dtype)


def pearsons_correlation(outputs, targets):
    """metrics.py:9-35, including its quirk (x target mean taken from the x outputs, :22).
    (bs, T, N, 2, D) -> x_corr, y_corr (bs, N, D), computed by asb_pearson (two passes over T in fp64); the
    per-sentence, length-masked form run_test needs is asb_eval_sentence_metrics."""
    if outputs.shape != targets.shape or outputs.dim() != 5 or outputs.shape[3] != 2:
        raise ValueError("expected two (bs, T, N_art, 2, N_samples) tensors of equal shape")
    dev, home = _lib.device_of(outputs, targets), outputs.device
    o = _as_f32_cuda(outputs, dev).contiguous()
    t = _as_f32_cuda(targets, dev).contiguous()
    B, T, N, _, D = o.shape
    x_corr = torch.empty(B, N, D, dtype=torch.float32, device=dev)
    y_corr = torch.empty(B, N, D, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        _lib.check(_lib.lib().asb_pearson(_lib.ptr(o), _lib.ptr(t), B, T, N, D, _lib.ptr(x_corr), _lib.ptr(y_corr),
                                          _lib.stream_ptr(dev)))
    dt = outputs.dtype if outputs.dtype.is_floating_point else torch.float32
    return x_corr.to(device=home, dtype=dt), y_corr.to(device=home, dtype=dt)


class P2CPDistance(nn.Module):
    """Validation metric, phoneme_to_articulation/encoder_decoder/metrics.py:7-26: mean over
    sentences of the length-masked mean P2CP in millimetres."""

    def __init__(self, dataset_config):
        super().__init__()
        self.dataset_config = dataset_config
        self.to_mm = self.dataset_config.RES * self.dataset_config.PIXEL_SPACING
        self.mean_p2cp = MeanP2CPDistance(reduction="none")

    def forward(self, outputs, targets, lengths):
        p2cp = self.mean_p2cp(outputs.transpose(-1, -2), targets.transpose(-1, -2)) * self.to_mm
        L = torch.as_tensor(lengths, device=p2cp.device).view(-1, 1)
        mask = (torch.arange(p2cp.shape[1], device=p2cp.device).view(1, -1) < L).unsqueeze(-1)
        per_sentence = (p2cp * mask).sum(dim=(1, 2)) / (L.view(-1) * p2cp.shape[2])
        return per_sentence.mean().cpu()
