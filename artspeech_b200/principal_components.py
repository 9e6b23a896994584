"""Drop-in ``PrincipalComponentsArtSpeech`` (phoneme_to_articulation/principal_components/models/rnn.py:36-109).

Same constructor (``vocab_size, indices_dict, embed_dim, hidden_size, rnn_dropout, rnn``), same
``forward(x, lengths)`` -> (bs, seq_len, num_components) and same ``state_dict`` keys
(``embedding``, ``rnn``, ``linear.0``, ``predictor.linear.{0,1,3,4,6,7}``).  The recurrent stack is
the one of ``ArtSpeech`` (``asb_model_forward_hidden``: tcgen05 recurrent kernels with W_hh in tensor
memory, GRU or LSTM); the tail - Linear+ReLU, the three LayerNorm/Linear stages of
``PrincipalComponentsPredictor`` and the tanh - is one fused kernel (``asb_pc_head_forward``).
The torch submodules are parameter containers only; there is no CPU or eager fallback.  Inference
only.
"""
from __future__ import annotations

import ctypes
from enum import Enum
from functools import reduce

import torch
import torch.nn as nn

from . import _lib
from .models import PRECISIONS, ArticulatorPredictor, _NativeHandle


class RNNType(Enum):
    """phoneme_to_articulation/__init__.py:47-49"""
    LSTM = nn.LSTM
    GRU = nn.GRU


def make_indices_dict(num_components):
    """helpers.py:94-114: {'a': 3, 'b': 2} -> {'a': [0, 1, 2], 'b': [3, 4]}"""
    indices_dict, start = {}, 0
    for key, val in num_components.items():
        indices_dict[key] = list(range(start, start + val))
        start += val
    return indices_dict


class PrincipalComponentsPredictor(nn.Module):
    """rnn.py:11-33 (parameter container; evaluated inside the fused head kernel)."""

    def __init__(self, in_features, num_components, hidden_features=256):
        super().__init__()
        self.linear = nn.Sequential(
            nn.LayerNorm([in_features]),
            nn.Linear(in_features=in_features, out_features=hidden_features),
            nn.ReLU(),
            nn.LayerNorm([hidden_features]),
            nn.Linear(in_features=hidden_features, out_features=hidden_features // 2),
            nn.ReLU(),
            nn.LayerNorm(hidden_features // 2),
            nn.Linear(in_features=hidden_features // 2, out_features=num_components),
        )

    def forward(self, inputs):
        raise RuntimeError("PrincipalComponentsPredictor is evaluated inside the fused CUDA head kernel; "
                           "it has no standalone forward in artspeech_b200")


class _HeadHandle:
    def __init__(self, ptr):
        self.ptr = ptr

    def __del__(self):
        try:
            if self.ptr:
                _lib.lib().asb_pc_head_destroy(self.ptr)
        except Exception:
            pass
        self.ptr = None


class PrincipalComponentsArtSpeech(nn.Module):
    """rnn.py:36-109.  ``precision``: "fast" (bf16 recurrent GEMM, tf32 elsewhere) or "fp32"; both for GRU and LSTM."""

    def __init__(self, vocab_size, indices_dict, embed_dim=64, hidden_size=128, rnn_dropout=0., rnn=RNNType.GRU,
                 precision="fast"):
        super().__init__()
        if isinstance(list(indices_dict.values())[0], int):
            indices_dict = make_indices_dict(indices_dict)
        self.latent_size = 1 + max(set(reduce(lambda l1, l2: l1 + l2, indices_dict.values())))
        self.embedding = nn.Embedding(vocab_size, embed_dim)
        if isinstance(rnn, str):
            rnn = RNNType[rnn.upper()]
        self._cell = _lib.CELL_LSTM if rnn.value is nn.LSTM else _lib.CELL_GRU
        self.rnn = rnn.value(embed_dim, hidden_size, num_layers=2, bidirectional=True, dropout=rnn_dropout,
                             batch_first=True)
        self.linear = nn.Sequential(nn.Linear(2 * hidden_size, hidden_size), nn.ReLU())
        self.predictor = PrincipalComponentsPredictor(in_features=hidden_size, num_components=self.latent_size,
                                                      hidden_features=256)
        if precision not in PRECISIONS:
            raise ValueError(f"precision must be one of {sorted(PRECISIONS)}")
        self.precision = precision
        self._dims = (vocab_size, embed_dim, hidden_size)
        self._native = None
        self._head = None
        self._keep = None
        self._workspace = None

    @property
    def total_parameters(self):
        return sum(p.numel() for p in self.parameters())

    def _apply(self, fn, *args, **kwargs):
        self._native = self._head = self._workspace = None
        return super()._apply(fn, *args, **kwargs)

    def _ensure_native(self, dev):
        key = (self.precision,) + tuple((p.data_ptr(), p._version) for p in self.parameters())
        if self._native is not None and self._native.key == key:
            return self._native.ptr, self._head.ptr
        self._native = self._head = None
        vocab, embed, hidden = self._dims
        sd = {k: v.detach() for k, v in self.named_parameters()}
        for k, v in sd.items():
            if not v.is_cuda or v.device != dev:
                raise RuntimeError(f"parameter {k} is on {v.device}; PrincipalComponentsArtSpeech runs on CUDA only "
                                   "(call .to('cuda')) - there is no CPU fallback")
            if v.dtype != torch.float32 or not v.is_contiguous():
                raise RuntimeError(f"parameter {k} must be contiguous float32")
        # recurrent stack: an asb_model whose decoder is never run (asb_model_forward_hidden); it still
        # wants one articulator's worth of decoder weights, which get zeros
        dummy = ArticulatorPredictor(hidden, 50).to(dev)
        for p in dummy.parameters():
            p.data.zero_()
        dsd = {k: v.detach() for k, v in dummy.named_parameters()}
        w = _lib.AsbWeights()
        w.embedding = sd["embedding.weight"].data_ptr()
        for layer in range(2):
            for d, sfx in enumerate(("", "_reverse")):
                w.rnn_w_ih[layer][d] = sd[f"rnn.weight_ih_l{layer}{sfx}"].data_ptr()
                w.rnn_w_hh[layer][d] = sd[f"rnn.weight_hh_l{layer}{sfx}"].data_ptr()
                w.rnn_b_ih[layer][d] = sd[f"rnn.bias_ih_l{layer}{sfx}"].data_ptr()
                w.rnn_b_hh[layer][d] = sd[f"rnn.bias_hh_l{layer}{sfx}"].data_ptr()
        w.linear_w = sd["linear.0.weight"].data_ptr()
        w.linear_b = sd["linear.0.bias"].data_ptr()
        keep = [dummy, dsd]

        def arr(name):
            a = (ctypes.c_void_p * 1)(dsd[name].data_ptr())
            keep.append(a)
            return ctypes.cast(a, ctypes.POINTER(ctypes.c_void_p))

        w.ln0_w, w.ln0_b = arr("linear.0.weight"), arr("linear.0.bias")
        w.fc1_w, w.fc1_b = arr("linear.1.weight"), arr("linear.1.bias")
        w.ln1_w, w.ln1_b = arr("linear.3.weight"), arr("linear.3.bias")
        w.fc2_w, w.fc2_b = arr("linear.4.weight"), arr("linear.4.bias")
        w.ln2_w, w.ln2_b = arr("linear.6.weight"), arr("linear.6.bias")
        w.xh_w, w.xh_b = arr("x_coords.weight"), arr("x_coords.bias")
        w.yh_w, w.yh_b = arr("y_coords.weight"), arr("y_coords.bias")
        pw = _lib.AsbPcWeights()
        pw.lin_w, pw.lin_b = sd["linear.0.weight"].data_ptr(), sd["linear.0.bias"].data_ptr()
        pw.ln0_w, pw.ln0_b = sd["predictor.linear.0.weight"].data_ptr(), sd["predictor.linear.0.bias"].data_ptr()
        pw.fc1_w, pw.fc1_b = sd["predictor.linear.1.weight"].data_ptr(), sd["predictor.linear.1.bias"].data_ptr()
        pw.ln1_w, pw.ln1_b = sd["predictor.linear.3.weight"].data_ptr(), sd["predictor.linear.3.bias"].data_ptr()
        pw.fc2_w, pw.fc2_b = sd["predictor.linear.4.weight"].data_ptr(), sd["predictor.linear.4.bias"].data_ptr()
        pw.ln2_w, pw.ln2_b = sd["predictor.linear.6.weight"].data_ptr(), sd["predictor.linear.6.bias"].data_ptr()
        pw.fc3_w, pw.fc3_b = sd["predictor.linear.7.weight"].data_ptr(), sd["predictor.linear.7.bias"].data_ptr()
        handle, head = ctypes.c_void_p(), ctypes.c_void_p()
        lib = _lib.lib()
        with torch.cuda.device(dev):
            _lib.check(lib.asb_model_create(ctypes.byref(w), vocab, 1, embed, hidden, 50, self._cell,
                                            PRECISIONS[self.precision], _lib.stream_ptr(dev), ctypes.byref(handle)))
            self._native = _NativeHandle(handle, key)
            _lib.check(lib.asb_pc_head_create(ctypes.byref(pw), hidden, self.latent_size, PRECISIONS[self.precision],
                                              _lib.stream_ptr(dev), ctypes.byref(head)))
            self._head = _HeadHandle(head)
        self._keep = keep
        return handle, head

    @torch.no_grad()
    def forward(self, x, lengths):
        """x (bs, seq) int64, lengths sorted descending -> (bs, max(lengths), num_components) float32."""
        if self.training and self.rnn.dropout > 0:
            raise NotImplementedError("inter-layer dropout (train mode with dropout > 0) is not implemented")
        dev = self.embedding.weight.device
        if dev.type != "cuda":
            raise RuntimeError("PrincipalComponentsArtSpeech runs on CUDA only (call .to('cuda')); no CPU fallback")
        if x.dim() != 2:
            raise ValueError(f"x must be (bs, seq_len), got {tuple(x.shape)}")
        B, T_in = x.shape
        L = [int(v) for v in (lengths.detach().cpu().tolist() if torch.is_tensor(lengths) else lengths)]
        if len(L) != B:
            raise RuntimeError(f"Expected `len(lengths)` to be equal to batch_size, but got {len(L)} (batch_size={B})")
        xd = x.detach().to(device=dev, dtype=torch.int64).contiguous()
        handle, head = self._ensure_native(dev)
        T = max(1, min(L[0], T_in)) if L else 0
        lib = _lib.lib()
        need = int(lib.asb_model_workspace_bytes(handle, B, T))
        if self._workspace is None or self._workspace.numel() < need or self._workspace.device != dev:
            self._workspace = torch.empty(need, dtype=torch.uint8, device=dev)
        hidden = torch.empty(B, T, 2 * self._dims[2], dtype=torch.float32, device=dev)
        out = torch.empty(B, T, self.latent_size, dtype=torch.float32, device=dev)
        c_len = (ctypes.c_int32 * B)(*L)
        with torch.cuda.device(dev):
            _lib.check(lib.asb_model_forward_hidden(handle, _lib.ptr(xd), T_in, c_len, None, B, _lib.ptr(hidden),
                                                    _lib.ptr(self._workspace), self._workspace.numel(),
                                                    _lib.stream_ptr(dev)))
            _lib.check(lib.asb_pc_head_forward(head, _lib.ptr(hidden), B * T, _lib.ptr(out), _lib.stream_ptr(dev)))
        return out
