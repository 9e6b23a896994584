"""Drop-in ``ArtSpeech`` for phoneme_to_articulation/encoder_decoder/models.py of the reference.

Same constructor, same ``forward(x, lengths)``, same parameter names and shapes (so reference
``state_dict`` checkpoints load unchanged, test_phoneme_to_articulation.py:70-72) - but the forward
pass runs entirely in the hand-written sm_100a kernels behind the C ABI (asb_model_*).

The torch submodules below are only PARAMETER CONTAINERS: they give identical initialisation
(same RNG consumption order as the reference constructor, models.py:109-120), identical
``state_dict`` keys, ``.to(device)`` and ``.parameters()``; the forward pass never calls their own
``forward`` and there is no CPU or eager fallback for it.  With autograd enabled (training,
train_phoneme_to_articulation.py:84-96) the module runs the unfused fp32 training path of
``artspeech_b200/training.py``: a forward that saves its activations and hand-written backward kernels
(GRU and LSTM cell, dropout included; no cuDNN / cuBLAS).
"""
from __future__ import annotations

import ctypes

import torch
import torch.nn as nn

from . import _lib

PRECISIONS = {"fast": _lib.PRECISION_FAST, "bf16": _lib.PRECISION_FAST, "fp32": _lib.PRECISION_FP32}


def _token_error(handle, vocab):
    """IndexError (what nn.Embedding raises, models.py:63,135) if a kernel of this model has seen a token
    id outside [0, vocab) since the last check.  Reads a mapped host word: no device synchronisation, so
    it reports the forwards that have COMPLETED (``check_tokens()`` synchronises first)."""
    bad, tok = ctypes.c_int32(0), ctypes.c_int64(0)
    _lib.check(_lib.lib().asb_model_token_status(handle, ctypes.byref(bad), ctypes.byref(tok), 1))
    if bad.value:
        raise IndexError(f"index out of range in self (token id {tok.value} is outside [0, {vocab}))")


def _host_token_check(x, vocab):
    """Tokens that arrive on the host are range-checked there, before anything is launched."""
    if not x.is_cuda and x.numel():
        lo, hi = int(x.min()), int(x.max())
        if lo < 0 or hi >= vocab:
            raise IndexError(f"index out of range in self (token id {lo if lo < 0 else hi} is outside [0, {vocab}))")


class _PredictorHandle:
    def __init__(self, ptr, key):
        self.ptr = ptr
        self.key = key

    def __del__(self):
        try:
            if self.ptr:
                _lib.lib().asb_predictor_destroy(self.ptr)
        except Exception:
            pass
        self.ptr = None


class _PredictorForward(torch.autograd.Function):
    """Standalone predictor under autograd: values from the kernel, gradient through the torch graph of
    the same parameters (library code: this standalone use is the transformer model's, outside the hot path)."""

    @staticmethod
    def forward(ctx, module, inputs, *params):
        ctx.module = module
        ctx.save_for_backward(inputs)
        with torch.no_grad():
            return module._native_forward(inputs)

    @staticmethod
    def backward(ctx, grad_out):
        module = ctx.module
        (inputs,) = ctx.saved_tensors
        params = list(module.parameters())
        with torch.enable_grad():
            xin = inputs.detach().requires_grad_(True)
            out = module._graph(xin)
            need = [xin] + [p for p in params if p.requires_grad]
            grads = torch.autograd.grad(out, need, grad_out.to(out.dtype), allow_unused=True)
        it = iter(grads[1:])
        return (None, grads[0]) + tuple(next(it) if p.requires_grad else None for p in params)


class ArticulatorPredictor(nn.Module):
    """models.py:7-33.  Inside ``ArtSpeech`` the predictors are parameter containers evaluated by the
    fused decoder kernel; called on its own (the transformer model builds its heads from this class,
    phoneme_to_articulation/transformer/models.py:6,337-339) ``forward`` runs ``asb_predictor_forward``:
    (..., in_features) -> (..., 2, n_samples), no sigmoid.  ``precision`` as for ``ArtSpeech``."""

    def __init__(self, in_features, n_samples, precision="fast"):
        super().__init__()
        self.precision = precision
        self._dims = (in_features, n_samples)
        self._native = None
        self.linear = nn.Sequential(
            nn.LayerNorm([in_features]),
            nn.Linear(in_features=in_features, out_features=256),
            nn.ReLU(),
            nn.LayerNorm([256]),
            nn.Linear(in_features=256, out_features=256),
            nn.ReLU(),
            nn.LayerNorm(256),
        )
        self.x_coords = nn.Linear(in_features=256, out_features=n_samples)
        self.y_coords = nn.Linear(in_features=256, out_features=n_samples)

    def _apply(self, fn, *args, **kwargs):
        self._native = None
        return super()._apply(fn, *args, **kwargs)

    def _ensure_native(self, dev):
        if self.precision not in PRECISIONS:
            raise ValueError(f"precision must be one of {sorted(PRECISIONS)}")
        key = (self.precision,) + tuple((p.data_ptr(), p._version) for p in self.parameters())
        if self._native is not None and self._native.key == key:
            return self._native.ptr
        self._native = None
        sd = {k: v.detach() for k, v in self.named_parameters()}
        for k, v in sd.items():
            if not v.is_cuda or v.device != dev:
                raise RuntimeError(f"parameter {k} is on {v.device}; artspeech_b200.ArticulatorPredictor runs on CUDA "
                                   "only (call .to('cuda')) - there is no CPU fallback")
            if v.dtype != torch.float32 or not v.is_contiguous():
                raise RuntimeError(f"parameter {k} must be contiguous float32")
        w = _lib.AsbPredictorWeights()
        for field, name in (("ln0_w", "linear.0.weight"), ("ln0_b", "linear.0.bias"), ("fc1_w", "linear.1.weight"),
                            ("fc1_b", "linear.1.bias"), ("ln1_w", "linear.3.weight"), ("ln1_b", "linear.3.bias"),
                            ("fc2_w", "linear.4.weight"), ("fc2_b", "linear.4.bias"), ("ln2_w", "linear.6.weight"),
                            ("ln2_b", "linear.6.bias"), ("xh_w", "x_coords.weight"), ("xh_b", "x_coords.bias"),
                            ("yh_w", "y_coords.weight"), ("yh_b", "y_coords.bias")):
            setattr(w, field, sd[name].data_ptr())
        handle = ctypes.c_void_p()
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().asb_predictor_create(ctypes.byref(w), self._dims[0], self._dims[1],
                                                       PRECISIONS[self.precision], _lib.stream_ptr(dev),
                                                       ctypes.byref(handle)))
        self._native = _PredictorHandle(handle, key)
        return handle

    def forward(self, inputs):
        """models.py:23-33: inputs (B, T, in_features) -> (B, T, 2, n_samples)."""
        if torch.is_grad_enabled() and (inputs.requires_grad or any(p.requires_grad for p in self.parameters())):
            return _PredictorForward.apply(self, inputs, *self.parameters())
        return self._native_forward(inputs)

    def _native_forward(self, inputs):
        dev = self.x_coords.weight.device
        if dev.type != "cuda":
            raise RuntimeError("artspeech_b200.ArticulatorPredictor runs on CUDA only (call .to('cuda')); no CPU fallback")
        in_features, n_samples = self._dims
        if inputs.shape[-1] != in_features:
            raise RuntimeError(f"expected inputs with last dimension {in_features}, got {tuple(inputs.shape)}")
        x = inputs.detach().to(device=dev, dtype=torch.float32).reshape(-1, in_features).contiguous()
        out = torch.empty(x.shape[0], 2, n_samples, dtype=torch.float32, device=dev)
        handle = self._ensure_native(dev)
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().asb_predictor_forward(handle, _lib.ptr(x), x.shape[0], _lib.ptr(out),
                                                        _lib.stream_ptr(dev)))
        if inputs.dim() == 3:                       # (B, T, C) -> (B, T, 2, D): the layout the kernel writes
            return out.view(*inputs.shape[:-1], 2, n_samples)
        # the reference stacks on dim 2 literally (models.py:31), whatever the input rank is
        lead = inputs.shape[:-1]
        return torch.stack([out[:, 0].reshape(*lead, n_samples), out[:, 1].reshape(*lead, n_samples)], dim=2)

    def _graph(self, inputs):
        """models.py:23-33 as a differentiable torch graph.  The forward pass never comes here (the
        predictors are evaluated inside the fused CUDA decoder); this graph exists only for the backward
        pass of the standalone predictor (``_PredictorForward``) and as the fp64 truth in the gradient tests."""
        linear_out = self.linear(inputs)
        return torch.stack([self.x_coords(linear_out), self.y_coords(linear_out)], dim=2)


class _NativeHandle:
    """Owns one asb_model (re-packed weights on one device) and frees it with the module."""

    def __init__(self, ptr, key):
        self.ptr = ptr
        self.key = key

    def __del__(self):
        try:
            if self.ptr:
                _lib.lib().asb_model_destroy(self.ptr)
        except Exception:
            pass
        self.ptr = None


class ArtSpeech(nn.Module):
    """models.py:99-145.  ``precision``: "fast" (fp16 GEMM operands on the tensor cores, fp32 accumulate/state;
    contours within 2e-3 of the fp32 reference) or "fp32" (operands split into hi + lo fp16 pairs, within 1e-5).
    ``rnn``: "gru" (the reference's cell, models.py:111) or "lstm" (the other member of RNNType,
    phoneme_to_articulation/__init__.py:47-49; same state_dict keys as nn.LSTM).
    Token ids outside [0, vocab_size) raise IndexError like nn.Embedding (models.py:135): at once for host
    tensors; for device tensors the kernels record it and the NEXT call or ``check_tokens()`` raises."""

    def __init__(self, vocab_size, n_articulators, embed_dim=64, hidden_size=128, n_samples=50, dropout=0.,
                 precision="fast", rnn="gru"):
        super().__init__()
        cells = {"gru": (nn.GRU, _lib.CELL_GRU), "lstm": (nn.LSTM, _lib.CELL_LSTM)}
        if str(rnn).lower() not in cells:
            raise ValueError(f"rnn must be one of {sorted(cells)}")
        rnn_cls, self._cell = cells[str(rnn).lower()]
        self.embedding = nn.Embedding(vocab_size, embed_dim)
        self.rnn = rnn_cls(embed_dim, hidden_size, num_layers=2, bidirectional=True, dropout=dropout,
                           batch_first=True)
        self.linear = nn.Sequential(nn.Linear(2 * hidden_size, hidden_size), nn.ReLU())
        self.predictors = nn.ModuleList([
            ArticulatorPredictor(hidden_size, n_samples) for _ in range(n_articulators)
        ])
        if precision not in PRECISIONS:
            raise ValueError(f"precision must be one of {sorted(PRECISIONS)}")
        self.precision = precision
        self._dims = (vocab_size, n_articulators, embed_dim, hidden_size, n_samples)
        self._native = None
        self._workspace = None

    @property
    def total_parameters(self):
        return sum(p.numel() for p in self.parameters())

    # ------------------------------------------------------------------ native handle
    def _weights_key(self):
        return (self.precision,) + tuple((p.data_ptr(), p._version) for p in self.parameters())

    def _ensure_native(self, dev):
        key = self._weights_key()
        if self._native is not None and self._native.key == key:
            return self._native.ptr
        self._native = None
        sd = {k: v.detach() for k, v in self.named_parameters()}
        for k, v in sd.items():
            if not v.is_cuda or v.device != dev:
                raise RuntimeError(f"parameter {k} is on {v.device}; artspeech_b200.ArtSpeech runs on CUDA only "
                                   "(call .to('cuda')) - there is no CPU fallback")
            if v.dtype != torch.float32 or not v.is_contiguous():
                raise RuntimeError(f"parameter {k} must be contiguous float32")
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
        n = self._dims[1]
        keep = []

        def arr(fmt):
            a = (ctypes.c_void_p * n)(*[sd[fmt.format(i)].data_ptr() for i in range(n)])
            keep.append(a)
            return ctypes.cast(a, ctypes.POINTER(ctypes.c_void_p))

        w.ln0_w, w.ln0_b = arr("predictors.{}.linear.0.weight"), arr("predictors.{}.linear.0.bias")
        w.fc1_w, w.fc1_b = arr("predictors.{}.linear.1.weight"), arr("predictors.{}.linear.1.bias")
        w.ln1_w, w.ln1_b = arr("predictors.{}.linear.3.weight"), arr("predictors.{}.linear.3.bias")
        w.fc2_w, w.fc2_b = arr("predictors.{}.linear.4.weight"), arr("predictors.{}.linear.4.bias")
        w.ln2_w, w.ln2_b = arr("predictors.{}.linear.6.weight"), arr("predictors.{}.linear.6.bias")
        w.xh_w, w.xh_b = arr("predictors.{}.x_coords.weight"), arr("predictors.{}.x_coords.bias")
        w.yh_w, w.yh_b = arr("predictors.{}.y_coords.weight"), arr("predictors.{}.y_coords.bias")
        vocab, n_art, embed, hidden, n_samples = self._dims
        handle = ctypes.c_void_p()
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().asb_model_create(
                ctypes.byref(w), vocab, n_art, embed, hidden, n_samples, self._cell,
                PRECISIONS[self.precision], _lib.stream_ptr(dev), ctypes.byref(handle)))
        self._native = _NativeHandle(handle, key)
        return handle

    def _apply(self, fn, *args, **kwargs):
        self._native = None  # parameters move / change dtype: re-pack lazily
        self._workspace = None
        return super()._apply(fn, *args, **kwargs)

    def invalidate_packed_weights(self):
        """Drop the packed copy of the weights so that the next forward re-packs them.  Needed only after
        edits that bypass autograd's version counter (``p.data.copy_()``, ``p.data.mul_()`` ...): ordinary
        in-place updates (optimizer steps, ``load_state_dict``, ``p.add_()``) are detected by themselves."""
        self._native = None

    def set_decoder(self, impl):
        """"tcgen05" (fp16 operands, default in fast precision), "tcgen05-tf32" (the round-1 kernel) or "mma"
        (mma.sync kernel) for the fused decoder."""
        dev = self.embedding.weight.device
        code = {"tcgen05": 2, "tcgen05-tf32": 1, "mma": 0}[impl]
        _lib.check(_lib.lib().asb_model_set_decoder(self._ensure_native(dev), code))

    def set_recurrent(self, impl):
        """"tcgen05" (default in fast precision: W_hh in tensor memory, tcgen05.mma with the gates behind
        tcgen05.ld) or "mma" (W_hh in registers, mma.sync; GRU only) for the two recurrent layers."""
        dev = self.embedding.weight.device
        code = {"tcgen05": 1, "mma": 0}[impl]
        _lib.check(_lib.lib().asb_model_set_recurrent(self._ensure_native(dev), code))

    def check_tokens(self):
        """Synchronise the model's device and raise IndexError if any forward since the last check was
        given a token id outside [0, vocab_size) (nn.Embedding semantics, models.py:135)."""
        if self._native is not None:
            torch.cuda.synchronize(self.embedding.weight.device)
            _token_error(self._native.ptr, self._dims[0])

    # ------------------------------------------------------------------ per-stage timing
    STAGES = ("gather", "gru_l0", "proj_l1", "gru_l1", "decoder")

    def profile(self, enable=True):
        """Record CUDA events around each kernel stage of subsequent forwards (asb_model_profile_*)."""
        dev = self.embedding.weight.device
        _lib.check(_lib.lib().asb_model_profile_enable(self._ensure_native(dev), 1 if enable else 0))

    def read_profile(self, reset=True):
        """{stage: (total_ms, calls)} accumulated since the last reset."""
        if self._native is None:
            return {}
        ms = (ctypes.c_float * 5)()
        calls = (ctypes.c_int32 * 5)()
        _lib.check(_lib.lib().asb_model_profile_read(self._native.ptr, ms, calls, 1 if reset else 0))
        return {name: (float(ms[k]), int(calls[k])) for k, name in enumerate(self.STAGES)}

    # ------------------------------------------------------------------ forward
    def forward(self, x, lengths, lengths_dev=None, workspace=None):
        """
        Args:
            x (torch.tensor): Torch tensor of shape (bs, seq_len), int64 tokens.
            lengths (list): Lengths of the input sequences sorted in decreasing order.
            lengths_dev (optional, extension): int32 CUDA tensor holding the same lengths; when given
                no host->device copy is enqueued, which makes the call capturable in a CUDA graph.
            workspace (optional, extension): uint8 CUDA tensor of at least ``workspace_bytes(bs, T)`` bytes
                used instead of the module's own scratch buffer (callers that run several forwards
                concurrently, or capture one in a CUDA graph, own their scratch).

        Return:
            output (torch.tensor): (bs, max(lengths), n_articulators, 2, n_samples) float32
        """
        needs_grad = torch.is_grad_enabled() and any(p.requires_grad for p in self.parameters())
        drops = self.training and self.rnn.dropout > 0
        if needs_grad or drops:
            if lengths_dev is not None or workspace is not None:
                raise ValueError("lengths_dev / workspace are inference-only extensions")
            L = lengths.detach().cpu().tolist() if torch.is_tensor(lengths) else list(lengths)
            L = [int(v) for v in L]
            # training (train_phoneme_to_articulation.py:84-96): forward with saved activations and
            # hand-written backward kernels (artspeech_b200/training.py), GRU and LSTM cell, dropout included
            from .training import ArtSpeechTrainingFunction
            return ArtSpeechTrainingFunction.apply(self, x, L, *self.parameters())
        return self._native_forward(x, lengths, lengths_dev, workspace)

    def workspace_bytes(self, batch, seq_len):
        """Scratch bytes one forward of a (batch, seq_len) input needs (asb_model_workspace_bytes)."""
        dev = self.embedding.weight.device
        return int(_lib.lib().asb_model_workspace_bytes(self._ensure_native(dev), int(batch), int(seq_len)))

    def _library_graph(self, x, lengths):
        """models.py:126-145 on the parameter containers as a differentiable torch graph (backward only)."""
        from torch.nn.utils.rnn import pack_padded_sequence, pad_packed_sequence
        packed = pack_padded_sequence(self.embedding(x), torch.as_tensor(lengths, dtype=torch.int64), batch_first=True)
        rnn_out, _ = pad_packed_sequence(self.rnn(packed)[0], batch_first=True)
        hidden = self.linear(rnn_out)
        return torch.sigmoid(torch.stack([predictor._graph(hidden) for predictor in self.predictors], dim=2))

    def _native_forward(self, x, lengths, lengths_dev=None, workspace=None):
        dev = self.embedding.weight.device
        if not dev.type == "cuda":
            raise RuntimeError("artspeech_b200.ArtSpeech runs on CUDA only (call .to('cuda')); no CPU fallback")
        if x.dim() != 2:
            raise ValueError(f"x must be (bs, seq_len), got {tuple(x.shape)}")
        B, T_in = x.shape
        if torch.is_tensor(lengths):
            L = [int(v) for v in lengths.detach().cpu().tolist()]
        else:
            L = [int(v) for v in lengths]
        if len(L) != B:
            raise RuntimeError(f"Expected `len(lengths)` to be equal to batch_size, but got {len(L)} "
                               f"(batch_size={B})")
        _host_token_check(x, self._dims[0])
        xd = x.detach().to(device=dev, dtype=torch.int64).contiguous()
        handle = self._ensure_native(dev)
        if not torch.cuda.is_current_stream_capturing():
            _token_error(handle, self._dims[0])      # lazily: ids seen by the forwards that have completed
        T = max(1, min(L[0], T_in)) if L else 0
        lib = _lib.lib()
        need = int(lib.asb_model_workspace_bytes(handle, B, T))
        if workspace is not None:
            if workspace.dtype != torch.uint8 or workspace.device != dev or workspace.numel() < need:
                raise ValueError(f"workspace must be a uint8 tensor of >= {need} bytes on {dev}")
        else:
            if self._workspace is None or self._workspace.numel() < need or self._workspace.device != dev:
                self._workspace = torch.empty(need, dtype=torch.uint8, device=dev)
            workspace = self._workspace
        vocab, n_art, _, _, n_samples = self._dims
        out = torch.empty(B, T, n_art, 2, n_samples, dtype=torch.float32, device=dev)
        c_len = (ctypes.c_int32 * B)(*L)
        ldev = None
        if lengths_dev is not None:
            if lengths_dev.dtype != torch.int32 or lengths_dev.device != dev or lengths_dev.numel() != B:
                raise ValueError("lengths_dev must be an int32 tensor of shape (bs,) on the model's device")
            ldev = _lib.ptr(lengths_dev)
        with torch.cuda.device(dev):
            _lib.check(lib.asb_model_forward_resident(handle, _lib.ptr(xd), T_in, c_len, ldev, B, _lib.ptr(out),
                                                      _lib.ptr(workspace), workspace.numel(),
                                                      _lib.stream_ptr(dev)))
        return out


class SimpleArtSpeech(nn.Module):
    """models.py:53-96 of the reference: Embedding -> Dropout -> Linear(embed_dim, hidden)+ReLU -> the same
    ``ArticulatorPredictor`` heads -> sigmoid; same constructor (note ``num_samples``), ``forward(x, lengths)``
    and state_dict keys.  There is no recurrence, so a frame's contours depend on its token only: the
    decoder kernel (``asb_model_decode``) is run once over the V embedding rows whenever the weights
    change, and ``forward`` is a row gather from that (V, N, 2, n_samples) table (``asb_gather_rows``),
    bound by the 4000 B per frame it writes.  ``lengths`` is accepted and ignored, as in the reference.
    Dropout is the identity in eval mode (models.py:64,92); train mode with dropout > 0 is not built."""

    def __init__(self, vocab_size, n_articulators, embed_dim=64, hidden_size=128, num_samples=50, dropout=0.,
                 precision="fast"):
        super().__init__()
        self.embedding = nn.Embedding(vocab_size, embed_dim)
        self.dropout = nn.Dropout(dropout)
        self.linear = nn.Sequential(nn.Linear(embed_dim, hidden_size), nn.ReLU())
        self.predictors = nn.ModuleList([
            ArticulatorPredictor(hidden_size, num_samples) for _ in range(n_articulators)
        ])
        if precision not in PRECISIONS:
            raise ValueError(f"precision must be one of {sorted(PRECISIONS)}")
        if embed_dim > 2 * hidden_size:
            raise ValueError("embed_dim must not exceed 2 * hidden_size (the decoder kernel's feature width)")
        self.precision = precision
        self._dims = (vocab_size, n_articulators, embed_dim, hidden_size, num_samples)
        self._native = None
        self._table = None
        self._keep = None

    @property
    def total_parameters(self):
        return sum(p.numel() for p in self.parameters())

    def _apply(self, fn, *args, **kwargs):
        self._native = None
        self._table = None
        return super()._apply(fn, *args, **kwargs)

    def _ensure_table(self, dev):
        key = (self.precision,) + tuple((p.data_ptr(), p._version) for p in self.parameters())
        if self._native is not None and self._native.key == key:
            return self._table
        self._native = None
        vocab, n_art, embed, hidden, n_samples = self._dims
        sd = {k: v.detach() for k, v in self.named_parameters()}
        for k, v in sd.items():
            if not v.is_cuda or v.device != dev:
                raise RuntimeError(f"parameter {k} is on {v.device}; artspeech_b200.SimpleArtSpeech runs on CUDA only "
                                   "(call .to('cuda')) - there is no CPU fallback")
            if v.dtype != torch.float32 or not v.is_contiguous():
                raise RuntimeError(f"parameter {k} must be contiguous float32")
        # The decoder kernel consumes (rows, 2*hidden) features through linear.0 (hidden, 2*hidden):
        # zero-pad the embedding rows and the Linear(embed_dim, hidden) weight to that width (layout
        # plumbing at pack time; the product with the padding is exactly zero).
        feat = torch.zeros(vocab, 2 * hidden, dtype=torch.float32, device=dev)
        feat[:, :embed] = sd["embedding.weight"]
        lin_w = torch.zeros(hidden, 2 * hidden, dtype=torch.float32, device=dev)
        lin_w[:, :embed] = sd["linear.0.weight"]
        # asb_model_create packs a recurrent stack as well; this model has none, so it gets zeros
        zeros = {(layer, what): torch.zeros(3 * hidden * (1 if what.startswith("b") else
                                                           (hidden if what == "w_hh" else (embed, 2 * hidden)[layer])),
                                            dtype=torch.float32, device=dev)
                 for layer in range(2) for what in ("w_ih", "w_hh", "b_ih", "b_hh")}
        w = _lib.AsbWeights()
        w.embedding = sd["embedding.weight"].data_ptr()
        for layer in range(2):
            for d in range(2):
                w.rnn_w_ih[layer][d] = zeros[(layer, "w_ih")].data_ptr()
                w.rnn_w_hh[layer][d] = zeros[(layer, "w_hh")].data_ptr()
                w.rnn_b_ih[layer][d] = zeros[(layer, "b_ih")].data_ptr()
                w.rnn_b_hh[layer][d] = zeros[(layer, "b_hh")].data_ptr()
        w.linear_w = lin_w.data_ptr()
        w.linear_b = sd["linear.0.bias"].data_ptr()
        keep = [feat, lin_w, zeros]

        def arr(fmt):
            a = (ctypes.c_void_p * n_art)(*[sd[fmt.format(i)].data_ptr() for i in range(n_art)])
            keep.append(a)
            return ctypes.cast(a, ctypes.POINTER(ctypes.c_void_p))

        w.ln0_w, w.ln0_b = arr("predictors.{}.linear.0.weight"), arr("predictors.{}.linear.0.bias")
        w.fc1_w, w.fc1_b = arr("predictors.{}.linear.1.weight"), arr("predictors.{}.linear.1.bias")
        w.ln1_w, w.ln1_b = arr("predictors.{}.linear.3.weight"), arr("predictors.{}.linear.3.bias")
        w.fc2_w, w.fc2_b = arr("predictors.{}.linear.4.weight"), arr("predictors.{}.linear.4.bias")
        w.ln2_w, w.ln2_b = arr("predictors.{}.linear.6.weight"), arr("predictors.{}.linear.6.bias")
        w.xh_w, w.xh_b = arr("predictors.{}.x_coords.weight"), arr("predictors.{}.x_coords.bias")
        w.yh_w, w.yh_b = arr("predictors.{}.y_coords.weight"), arr("predictors.{}.y_coords.bias")
        handle = ctypes.c_void_p()
        table = torch.empty(vocab, n_art, 2, n_samples, dtype=torch.float32, device=dev)
        lib = _lib.lib()
        with torch.cuda.device(dev):
            _lib.check(lib.asb_model_create(ctypes.byref(w), vocab, n_art, embed, hidden, n_samples, _lib.CELL_GRU,
                                            PRECISIONS[self.precision], _lib.stream_ptr(dev), ctypes.byref(handle)))
            self._native = _NativeHandle(handle, key)
            _lib.check(lib.asb_model_decode(handle, _lib.ptr(feat), vocab, _lib.ptr(table), _lib.stream_ptr(dev)))
        self._table = table
        self._keep = keep
        return table

    def forward(self, x, lengths=None):
        """x (bs, seq_len) int64 -> (bs, seq_len, n_articulators, 2, n_samples) float32 (models.py:84-96).
        Under autograd (or in train mode with dropout) it runs the native training path
        (``training.SimpleArtSpeechTrainingFunction``)."""
        needs_grad = torch.is_grad_enabled() and any(p.requires_grad for p in self.parameters())
        if needs_grad or (self.training and self.dropout.p > 0):
            from .training import SimpleArtSpeechTrainingFunction
            return SimpleArtSpeechTrainingFunction.apply(self, x, lengths, *self.parameters())
        return self._native_forward(x, lengths)

    def _library_graph(self, x, lengths=None):
        """models.py:84-96 on the parameter containers as a differentiable torch graph (backward only)."""
        hidden = self.linear(self.dropout(self.embedding(x)))
        return torch.sigmoid(torch.stack([predictor._graph(hidden) for predictor in self.predictors], dim=2))

    def _native_forward(self, x, lengths=None):
        dev = self.embedding.weight.device
        if dev.type != "cuda":
            raise RuntimeError("artspeech_b200.SimpleArtSpeech runs on CUDA only (call .to('cuda')); no CPU fallback")
        if x.dim() != 2:
            raise ValueError(f"x must be (bs, seq_len), got {tuple(x.shape)}")
        table = self._ensure_table(dev)
        vocab, n_art, _, _, n_samples = self._dims
        row = n_art * 2 * n_samples
        if row % 4:
            raise ValueError("n_articulators * 2 * num_samples must be a multiple of 4")
        _host_token_check(x, vocab)
        xd = x.detach().to(device=dev, dtype=torch.int64).contiguous()
        B, T = xd.shape
        out = torch.empty(B, T, n_art, 2, n_samples, dtype=torch.float32, device=dev)
        lib = _lib.lib()
        if not torch.cuda.is_current_stream_capturing():
            _token_error(self._native.ptr, vocab)
        with torch.cuda.device(dev):
            _lib.check(lib.asb_gather_rows_flagged(_lib.ptr(table), _lib.ptr(xd), B * T, vocab, row, _lib.ptr(out),
                                                   ctypes.c_void_p(lib.asb_model_token_flag_ptr(self._native.ptr)),
                                                   _lib.stream_ptr(dev)))
        return out

    def check_tokens(self):
        """Synchronise and raise IndexError if a token id outside [0, vocab_size) was seen (models.py:63)."""
        if self._native is not None:
            torch.cuda.synchronize(self.embedding.weight.device)
            _token_error(self._native.ptr, self._dims[0])
