"""Training path of the drop-in models: forward with saved activations and hand-written backward.

``train_phoneme_to_articulation.py:84-96`` runs ``loss.backward()`` through ``ArtSpeech``.  The fused
inference kernels save no activations, so under autograd the modules run the unfused fp32 path below,
assembled from the primitives of ``csrc/train_ops.cu`` (``asb_gemm``, ``asb_layernorm_*``,
``asb_gru_train_*`` ...): every forward and backward kernel is this library's own - no cuDNN / cuBLAS
kernel is launched - and the weights are read straight from the ``nn.Parameter`` storage (nothing is
re-packed after an optimizer step).  Gradients equal the reference model's autograd gradients
(``tests/golden/model_grad.npz``).  Inter-layer GRU dropout (``models.py:111``) and ``nn.Dropout``
(``models.py:65``) are applied in train mode with a counter-based generator seeded from torch's RNG.
"""
from __future__ import annotations

import ctypes

import torch

from . import _lib

_F32 = torch.float32


def _st(dev):
    return _lib.stream_ptr(dev)


def _check(t, name):
    if not t.is_cuda or t.dtype != _F32:
        raise RuntimeError(f"{name} must be a float32 CUDA tensor (artspeech_b200 has no CPU fallback)")


def _ld(t):
    """Row stride of a 2-D view whose last dimension is contiguous."""
    assert t.dim() == 2 and t.stride(1) == 1, (t.shape, t.stride())
    return t.stride(0)


def linear_forward(x, w, b, out, act=0):
    """out = act(x W^T + b).  x (M, K), w (N, K), out (M, N) views with contiguous last dimension."""
    M, K = x.shape
    N = w.shape[0]
    _lib.check(_lib.lib().asb_gemm(_lib.ptr(x), _lib.ptr(w), _lib.ptr(out), M, N, K, _ld(x), 1, 1, _ld(w), _ld(out),
                                   _lib.ptr(b) if b is not None else None, act, 0, _st(x.device)))
    return out


def linear_backward_input(dy, w, dx, accumulate=False):
    """dx (+)= dy W.  dy (M, N), w (N, K), dx (M, K)."""
    M, N = dy.shape
    K = w.shape[1]
    _lib.check(_lib.lib().asb_gemm(_lib.ptr(dy), _lib.ptr(w), _lib.ptr(dx), M, K, N, _ld(dy), 1, _ld(w), 1, _ld(dx),
                                   None, 0, 1 if accumulate else 0, _st(dy.device)))
    return dx


def linear_backward_weight(dy, x, dw, db=None):
    """dW = dy^T x, db = column sums of dy.  dy (M, N), x (M, K), dw (N, K) contiguous."""
    M, N = dy.shape
    K = x.shape[1]
    lib = _lib.lib()
    _lib.check(lib.asb_gemm(_lib.ptr(dy), _lib.ptr(x), _lib.ptr(dw), N, K, M, 1, _ld(dy), _ld(x), 1, K, None, 0, 0,
                            _st(dy.device)))
    if db is not None:
        _lib.check(lib.asb_colsum(_lib.ptr(dy), M, N, _ld(dy), _lib.ptr(db), 0, _st(dy.device)))
    return dw


def layernorm_forward(x, gamma, beta):
    M, C = x.shape
    xhat, y = torch.empty_like(x), torch.empty_like(x)
    rstd = torch.empty(M, dtype=_F32, device=x.device)
    _lib.check(_lib.lib().asb_layernorm_fwd(_lib.ptr(x), M, C, _lib.ptr(gamma), _lib.ptr(beta), _lib.ptr(xhat),
                                            _lib.ptr(rstd), _lib.ptr(y), _st(x.device)))
    return xhat, rstd, y


def layernorm_backward(dy, xhat, rstd, gamma, relu_src, dx, accumulate_dx, dgamma, dbeta):
    M, C = dy.shape
    _lib.check(_lib.lib().asb_layernorm_bwd(
        _lib.ptr(dy), _lib.ptr(xhat), _lib.ptr(rstd), _lib.ptr(gamma), _lib.ptr(relu_src) if relu_src is not None else None,
        M, C, _lib.ptr(dx), 1 if accumulate_dx else 0, _lib.ptr(dgamma), _lib.ptr(dbeta), 0, _st(dy.device)))
    return dx


def dropout(x, p, seed):
    out = torch.empty_like(x)
    _lib.check(_lib.lib().asb_dropout(_lib.ptr(x), _lib.ptr(out), x.numel(), float(p), int(seed), _st(x.device)))
    return out


def _fresh_seed():
    """A 63-bit seed drawn from torch's default CPU generator: torch.manual_seed makes training reproducible."""
    return int(torch.randint(0, 2 ** 62, (1,), dtype=torch.int64).item())


# ------------------------------------------------------------------------------------------ predictors
def predictors_forward(hidden, predictors, out):
    """models.py:7-33 x N + stack + sigmoid (models.py:141-145).  hidden (M, H); out (M, N, 2, D) receives the
    contours; returns the per-articulator saved activations."""
    M = hidden.shape[0]
    n = len(predictors)
    d = predictors[0].x_coords.weight.shape[0]
    flat = out.view(M, n * 2 * d)
    saved = []
    for a, pr in enumerate(predictors):
        ln0, fc1, ln1, fc2, ln2 = pr.linear[0], pr.linear[1], pr.linear[3], pr.linear[4], pr.linear[6]
        xh0, rs0, y0 = layernorm_forward(hidden, ln0.weight, ln0.bias)
        f1 = linear_forward(y0, fc1.weight, fc1.bias, torch.empty(M, fc1.weight.shape[0], dtype=_F32, device=hidden.device), act=1)
        xh1, rs1, y1 = layernorm_forward(f1, ln1.weight, ln1.bias)
        f2 = linear_forward(y1, fc2.weight, fc2.bias, torch.empty(M, fc2.weight.shape[0], dtype=_F32, device=hidden.device), act=1)
        xh2, rs2, y2 = layernorm_forward(f2, ln2.weight, ln2.bias)
        base = a * 2 * d
        linear_forward(y2, pr.x_coords.weight, pr.x_coords.bias, flat[:, base:base + d], act=2)
        linear_forward(y2, pr.y_coords.weight, pr.y_coords.bias, flat[:, base + d:base + 2 * d], act=2)
        saved.append((xh0, rs0, y0, f1, xh1, rs1, y1, f2, xh2, rs2, y2))
    return saved


def predictors_backward(dz, predictors, saved, d_hidden, grads):
    """dz (M, N*2*D) = gradient wrt the pre-sigmoid head outputs.  Adds the gradient wrt ``hidden`` into d_hidden
    (M, H) and stores every predictor parameter's gradient in ``grads`` (id(param) -> tensor)."""
    M = dz.shape[0]
    dev = dz.device
    n = len(predictors)
    d = predictors[0].x_coords.weight.shape[0]
    new = lambda p: torch.empty_like(p)
    for a, pr in enumerate(predictors):
        ln0, fc1, ln1, fc2, ln2 = pr.linear[0], pr.linear[1], pr.linear[3], pr.linear[4], pr.linear[6]
        xh0, rs0, y0, f1, xh1, rs1, y1, f2, xh2, rs2, y2 = saved[a]
        base = a * 2 * d
        dzx, dzy = dz[:, base:base + d], dz[:, base + d:base + 2 * d]
        for head, g in ((pr.x_coords, dzx), (pr.y_coords, dzy)):
            grads[id(head.weight)] = linear_backward_weight(g, y2, new(head.weight), None)
            gb = new(head.bias)
            _lib.check(_lib.lib().asb_colsum(_lib.ptr(g), M, d, _ld(g), _lib.ptr(gb), 0, _st(dev)))
            grads[id(head.bias)] = gb
        d_a2 = torch.empty(M, y2.shape[1], dtype=_F32, device=dev)
        linear_backward_input(dzx, pr.x_coords.weight, d_a2)
        linear_backward_input(dzy, pr.y_coords.weight, d_a2, accumulate=True)
        # LayerNorm 2 and the ReLU in front of it
        g2, b2 = new(ln2.weight), new(ln2.bias)
        d_pre2 = layernorm_backward(d_a2, xh2, rs2, ln2.weight, f2, torch.empty_like(d_a2), False, g2, b2)
        grads[id(ln2.weight)], grads[id(ln2.bias)] = g2, b2
        gw, gb = new(fc2.weight), new(fc2.bias)
        linear_backward_weight(d_pre2, y1, gw, gb)
        grads[id(fc2.weight)], grads[id(fc2.bias)] = gw, gb
        d_a1 = linear_backward_input(d_pre2, fc2.weight, torch.empty(M, y1.shape[1], dtype=_F32, device=dev))
        g1, b1 = new(ln1.weight), new(ln1.bias)
        d_pre1 = layernorm_backward(d_a1, xh1, rs1, ln1.weight, f1, torch.empty_like(d_a1), False, g1, b1)
        grads[id(ln1.weight)], grads[id(ln1.bias)] = g1, b1
        gw, gb = new(fc1.weight), new(fc1.bias)
        linear_backward_weight(d_pre1, y0, gw, gb)
        grads[id(fc1.weight)], grads[id(fc1.bias)] = gw, gb
        d_a0 = linear_backward_input(d_pre1, fc1.weight, torch.empty(M, y0.shape[1], dtype=_F32, device=dev))
        g0, b0 = new(ln0.weight), new(ln0.bias)
        layernorm_backward(d_a0, xh0, rs0, ln0.weight, None, d_hidden, True, g0, b0)
        grads[id(ln0.weight)], grads[id(ln0.bias)] = g0, b0


# ------------------------------------------------------------------------------------------ GRU layer
def _gru_params(rnn, layer):
    out = []
    for sfx in ("", "_reverse"):
        out.append(tuple(getattr(rnn, f"{name}_l{layer}{sfx}") for name in ("weight_ih", "weight_hh", "bias_ih", "bias_hh")))
    return out


def rnn_layer_forward(x, rnn, layer, lengths_dev, B, T):
    """One bidirectional nn.GRU / nn.LSTM layer: x (M, in) -> out (M, 2H) with packed-sequence semantics;
    returns (out, gates)."""
    M = x.shape[0]
    H = rnn.hidden_size
    G = 4 if isinstance(rnn, torch.nn.LSTM) else 3
    dev = x.device
    (wih_f, whh_f, bih_f, bhh_f), (wih_b, whh_b, bih_b, bhh_b) = _gru_params(rnn, layer)
    gi = torch.empty(M, 2, G * H, dtype=_F32, device=dev)
    flat = gi.view(M, 2 * G * H)
    linear_forward(x, wih_f, bih_f, flat[:, :G * H])
    linear_forward(x, wih_b, bih_b, flat[:, G * H:])
    out = torch.zeros(M, 2 * H, dtype=_F32, device=dev)
    gates = torch.empty(M, 2, G + 1, H, dtype=_F32, device=dev)
    fn = _lib.lib().asb_lstm_train_forward if G == 4 else _lib.lib().asb_gru_train_forward
    _lib.check(fn(_lib.ptr(whh_f), _lib.ptr(whh_b), _lib.ptr(bhh_f), _lib.ptr(bhh_b), _lib.ptr(gi), _lib.ptr(lengths_dev),
                  B, T, H, _lib.ptr(out), _lib.ptr(gates), _st(dev)))
    return out, gates


def rnn_layer_backward(d_out, x, out, gates, rnn, layer, lengths_dev, B, T, grads, need_dx=True):
    """BPTT through one bidirectional layer.  Returns the gradient wrt its input x (M, in) (or None)."""
    M = x.shape[0]
    H = rnn.hidden_size
    G = 4 if isinstance(rnn, torch.nn.LSTM) else 3
    dev = x.device
    lib = _lib.lib()
    params = _gru_params(rnn, layer)
    dgi = torch.zeros(M, 2, G * H, dtype=_F32, device=dev)
    hprev = torch.zeros(M, 2, H, dtype=_F32, device=dev)
    if G == 4:
        dgh = dgi            # every LSTM gate pre-activation is gi + W_hh h + b_hh
        _lib.check(lib.asb_lstm_train_backward(_lib.ptr(params[0][1]), _lib.ptr(params[1][1]), _lib.ptr(d_out),
                                               _lib.ptr(out), _lib.ptr(gates), _lib.ptr(lengths_dev), B, T, H,
                                               _lib.ptr(dgi), _lib.ptr(hprev), _st(dev)))
    else:
        dgh = torch.zeros(M, 2, G * H, dtype=_F32, device=dev)
        _lib.check(lib.asb_gru_train_backward(_lib.ptr(params[0][1]), _lib.ptr(params[1][1]), _lib.ptr(d_out),
                                              _lib.ptr(out), _lib.ptr(gates), _lib.ptr(lengths_dev), B, T, H,
                                              _lib.ptr(dgi), _lib.ptr(dgh), _lib.ptr(hprev), _st(dev)))
    dgi2, dgh2, hp2 = dgi.view(M, 2 * G * H), dgh.view(M, 2 * G * H), hprev.view(M, 2 * H)
    dx = torch.empty_like(x) if need_dx else None
    for d, (wih, whh, bih, bhh) in enumerate(params):
        gi_d, gh_d, hp_d = dgi2[:, d * G * H:(d + 1) * G * H], dgh2[:, d * G * H:(d + 1) * G * H], hp2[:, d * H:(d + 1) * H]
        grads[id(wih)] = linear_backward_weight(gi_d, x, torch.empty_like(wih), None)
        grads[id(whh)] = linear_backward_weight(gh_d, hp_d, torch.empty_like(whh), None)
        gb_i, gb_h = torch.empty_like(bih), torch.empty_like(bhh)
        _lib.check(lib.asb_colsum(_lib.ptr(gi_d), M, G * H, _ld(gi_d), _lib.ptr(gb_i), 0, _st(dev)))
        _lib.check(lib.asb_colsum(_lib.ptr(gh_d), M, G * H, _ld(gh_d), _lib.ptr(gb_h), 0, _st(dev)))
        grads[id(bih)], grads[id(bhh)] = gb_i, gb_h
        if need_dx:
            linear_backward_input(gi_d, wih, dx, accumulate=d > 0)
    return dx


# ------------------------------------------------------------------------------------------ whole models
def _embedding_forward(weight, tokens):
    rows = tokens.numel()
    out = torch.empty(rows, weight.shape[1], dtype=_F32, device=weight.device)
    _lib.check(_lib.lib().asb_gather_rows(_lib.ptr(weight), _lib.ptr(tokens), rows, weight.shape[0], weight.shape[1],
                                          _lib.ptr(out), _st(weight.device)))
    return out


def _embedding_backward(d_emb, tokens, weight):
    g = torch.zeros_like(weight)
    _lib.check(_lib.lib().asb_scatter_add_rows(_lib.ptr(d_emb), _lib.ptr(tokens), tokens.numel(), weight.shape[1],
                                               weight.shape[0], _lib.ptr(g), _st(weight.device)))
    return g


class ArtSpeechTrainingFunction(torch.autograd.Function):
    """``ArtSpeech.forward`` (models.py:126-145) under autograd: native forward that saves its activations,
    native backward (BPTT through both GRU layers, the predictors, the embedding)."""

    @staticmethod
    def forward(ctx, module, x, lengths, *params):
        from .models import _host_token_check
        dev = module.embedding.weight.device
        if dev.type != "cuda":
            raise RuntimeError("artspeech_b200.ArtSpeech runs on CUDA only (call .to('cuda')); no CPU fallback")
        for p in params:
            _check(p, "every parameter")
        L = [int(v) for v in lengths]
        B, T_in = x.shape
        if len(L) != B:
            raise RuntimeError(f"Expected `len(lengths)` to be equal to batch_size, but got {len(L)} (batch_size={B})")
        if any(v <= 0 for v in L):
            raise RuntimeError("Length of all samples has to be greater than 0, but found an element in 'lengths' that is <= 0")
        if any(a < b for a, b in zip(L, L[1:])):
            raise RuntimeError("`lengths` array must be sorted in decreasing order when `enforce_sorted` is True.")
        T = min(L[0], T_in)
        _host_token_check(x, module._dims[0])
        tok = x.detach().to(device=dev, dtype=torch.int64)[:, :T].contiguous()
        if tok.numel() and (int(tok.min()) < 0 or int(tok.max()) >= module._dims[0]):
            raise IndexError("index out of range in self")
        ldev = torch.tensor([min(v, T) for v in L], dtype=torch.int32, device=dev)
        M = B * T
        n_art, d = module._dims[1], module._dims[4]
        with torch.cuda.device(dev):
            emb = _embedding_forward(module.embedding.weight, tok)
            out0, gates0 = rnn_layer_forward(emb, module.rnn, 0, ldev, B, T)
            p_drop = float(module.rnn.dropout) if module.training else 0.0
            seed = _fresh_seed() if p_drop > 0 else 0
            in1 = dropout(out0, p_drop, seed) if p_drop > 0 else out0
            out1, gates1 = rnn_layer_forward(in1, module.rnn, 1, ldev, B, T)
            lin = module.linear[0]
            hidden = linear_forward(out1, lin.weight, lin.bias, torch.empty(M, lin.weight.shape[0], dtype=_F32, device=dev), act=1)
            y = torch.empty(B, T, n_art, 2, d, dtype=_F32, device=dev)
            saved = predictors_forward(hidden, module.predictors, y.view(M, n_art, 2, d))
        ctx.module, ctx.saved = module, (tok, ldev, emb, out0, gates0, in1, out1, gates1, hidden, saved, y, p_drop, seed, B, T)
        ctx.n_params = len(params)
        return y

    @staticmethod
    def backward(ctx, grad_out):
        module = ctx.module
        tok, ldev, emb, out0, gates0, in1, out1, gates1, hidden, saved, y, p_drop, seed, B, T = ctx.saved
        dev = y.device
        M = B * T
        grads = {}
        with torch.cuda.device(dev):
            dy = grad_out.to(_F32).contiguous()
            dz = torch.empty(M, y.numel() // M, dtype=_F32, device=dev)
            _lib.check(_lib.lib().asb_sigmoid_bwd(_lib.ptr(dy), _lib.ptr(y), _lib.ptr(dz), y.numel(), _st(dev)))
            d_hidden = torch.zeros_like(hidden)
            predictors_backward(dz, module.predictors, saved, d_hidden, grads)
            _lib.check(_lib.lib().asb_relu_mask(_lib.ptr(d_hidden), _lib.ptr(hidden), hidden.numel(), _st(dev)))
            lin = module.linear[0]
            gw, gb = torch.empty_like(lin.weight), torch.empty_like(lin.bias)
            linear_backward_weight(d_hidden, out1, gw, gb)
            grads[id(lin.weight)], grads[id(lin.bias)] = gw, gb
            d_out1 = linear_backward_input(d_hidden, lin.weight, torch.empty_like(out1))
            d_in1 = rnn_layer_backward(d_out1, in1, out1, gates1, module.rnn, 1, ldev, B, T, grads)
            d_out0 = dropout(d_in1, p_drop, seed) if p_drop > 0 else d_in1
            d_emb = rnn_layer_backward(d_out0, emb, out0, gates0, module.rnn, 0, ldev, B, T, grads)
            grads[id(module.embedding.weight)] = _embedding_backward(d_emb, tok, module.embedding.weight)
        out = tuple(grads.get(id(p)) if p.requires_grad else None for p in module.parameters())
        return (None, None, None) + out


class SimpleArtSpeechTrainingFunction(torch.autograd.Function):
    """``SimpleArtSpeech.forward`` (models.py:84-96) under autograd: embedding -> dropout -> Linear + ReLU ->
    predictors -> sigmoid, forward with saved activations and native backward."""

    @staticmethod
    def forward(ctx, module, x, lengths, *params):
        from .models import _host_token_check
        dev = module.embedding.weight.device
        if dev.type != "cuda":
            raise RuntimeError("artspeech_b200.SimpleArtSpeech runs on CUDA only (call .to('cuda')); no CPU fallback")
        for p in params:
            _check(p, "every parameter")
        _host_token_check(x, module._dims[0])
        tok = x.detach().to(device=dev, dtype=torch.int64).contiguous()
        if tok.numel() and (int(tok.min()) < 0 or int(tok.max()) >= module._dims[0]):
            raise IndexError("index out of range in self")
        B, T = tok.shape
        M = B * T
        n_art, d = module._dims[1], module._dims[4]
        with torch.cuda.device(dev):
            emb = _embedding_forward(module.embedding.weight, tok)
            p_drop = float(module.dropout.p) if module.training else 0.0
            seed = _fresh_seed() if p_drop > 0 else 0
            feat = dropout(emb, p_drop, seed) if p_drop > 0 else emb
            lin = module.linear[0]
            hidden = linear_forward(feat, lin.weight, lin.bias, torch.empty(M, lin.weight.shape[0], dtype=_F32, device=dev), act=1)
            y = torch.empty(B, T, n_art, 2, d, dtype=_F32, device=dev)
            saved = predictors_forward(hidden, module.predictors, y.view(M, n_art, 2, d))
        ctx.module, ctx.saved = module, (tok, feat, hidden, saved, y, p_drop, seed)
        return y

    @staticmethod
    def backward(ctx, grad_out):
        module = ctx.module
        tok, feat, hidden, saved, y, p_drop, seed = ctx.saved
        dev = y.device
        M = hidden.shape[0]
        grads = {}
        with torch.cuda.device(dev):
            dy = grad_out.to(_F32).contiguous()
            dz = torch.empty(M, y.numel() // M, dtype=_F32, device=dev)
            _lib.check(_lib.lib().asb_sigmoid_bwd(_lib.ptr(dy), _lib.ptr(y), _lib.ptr(dz), y.numel(), _st(dev)))
            d_hidden = torch.zeros_like(hidden)
            predictors_backward(dz, module.predictors, saved, d_hidden, grads)
            _lib.check(_lib.lib().asb_relu_mask(_lib.ptr(d_hidden), _lib.ptr(hidden), hidden.numel(), _st(dev)))
            lin = module.linear[0]
            gw, gb = torch.empty_like(lin.weight), torch.empty_like(lin.bias)
            linear_backward_weight(d_hidden, feat, gw, gb)
            grads[id(lin.weight)], grads[id(lin.bias)] = gw, gb
            d_feat = linear_backward_input(d_hidden, lin.weight, torch.empty_like(feat))
            d_emb = dropout(d_feat, p_drop, seed) if p_drop > 0 else d_feat
            grads[id(module.embedding.weight)] = _embedding_backward(d_emb, tok, module.embedding.weight)
        out = tuple(grads.get(id(p)) if p.requires_grad else None for p in module.parameters())
        return (None, None, None) + out
