"""Oracle restatement of the encoder_decoder ``ArtSpeech`` forward.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

Follows ``phoneme_to_articulation/encoder_decoder/models.py``:
  * ArticulatorPredictor           :7-33
  * ArtSpeech.__init__ / forward   :99-145

The forward is written functionally over a ``state_dict`` (the reference's key
names) so that it can run in fp32 (the tolerance target) or fp64
(truth) without building the reference classes.  Two recurrences are offered:
``rnn="stock"`` drives ``torch.nn.GRU`` + pack/pad exactly like models.py:136-138;
``rnn="equations"`` evaluates the GRU cell equations row by row (PyTorch GRU
documentation; gate order r, z, n) and is used to cross-check the first.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.nn.utils.rnn import pack_padded_sequence, pad_packed_sequence

LN_EPS = 1e-5  # nn.LayerNorm default, models.py:11,14,17


def make_state_dict(vocab_size, n_articulators, embed_dim=64, hidden_size=128,
                    n_samples=50, seed=0, cell="gru"):
    """Random-init weights with the reference's parameter names and the same
    RNG consumption order as ``ArtSpeech.__init__`` (models.py:109-120), so that
    ``torch.manual_seed(seed)`` reproduces the reference's initial weights
    bit for bit (checked in tests/test_oracle_pinning.py against golden sums).
    ``cell="lstm"`` builds the 4-gate variant (RNNType.LSTM,
    phoneme_to_articulation/__init__.py:47-49).
    """
    torch.manual_seed(seed)
    sd = {}
    emb = nn.Embedding(vocab_size, embed_dim)
    rnn_cls = nn.GRU if cell == "gru" else nn.LSTM
    rnn = rnn_cls(embed_dim, hidden_size, num_layers=2, bidirectional=True, batch_first=True)
    lin = nn.Linear(2 * hidden_size, hidden_size)
    sd["embedding.weight"] = emb.weight.detach().clone()
    for k, v in rnn.state_dict().items():
        sd[f"rnn.{k}"] = v.detach().clone()
    sd["linear.0.weight"] = lin.weight.detach().clone()
    sd["linear.0.bias"] = lin.bias.detach().clone()
    for i in range(n_articulators):
        ln0 = nn.LayerNorm([hidden_size])
        fc1 = nn.Linear(hidden_size, 256)
        ln1 = nn.LayerNorm([256])
        fc2 = nn.Linear(256, 256)
        ln2 = nn.LayerNorm(256)
        xh = nn.Linear(256, n_samples)
        yh = nn.Linear(256, n_samples)
        p = f"predictors.{i}."
        for idx, mod in ((0, ln0), (1, fc1), (3, ln1), (4, fc2), (6, ln2)):
            sd[p + f"linear.{idx}.weight"] = mod.weight.detach().clone()
            sd[p + f"linear.{idx}.bias"] = mod.bias.detach().clone()
        sd[p + "x_coords.weight"] = xh.weight.detach().clone()
        sd[p + "x_coords.bias"] = xh.bias.detach().clone()
        sd[p + "y_coords.weight"] = yh.weight.detach().clone()
        sd[p + "y_coords.bias"] = yh.bias.detach().clone()
    return sd


def randomize_affine(sd, seed=1):
    """Perturb LayerNorm affine terms and biases (which init to 1 / 0) so that
    parity tests exercise them.  Returns a new dict."""
    g = torch.Generator().manual_seed(seed)
    out = {}
    for k, v in sd.items():
        v = v.clone()
        if ".linear.0." in k or ".linear.3." in k or ".linear.6." in k:
            if k.startswith("predictors"):
                v = v + 0.1 * torch.randn(v.shape, generator=g, dtype=v.dtype)
        out[k] = v
    return out


def n_articulators_of(sd):
    n = 0
    while f"predictors.{n}.x_coords.weight" in sd:
        n += 1
    return n


def _cell_of(sd):
    gates = sd["rnn.weight_hh_l0"].shape[0] // sd["rnn.weight_hh_l0"].shape[1]
    return {3: "gru", 4: "lstm"}[gates]


def _rnn_stock(sd, embed, lengths, dtype):
    """models.py:136-138 (pack -> nn.GRU -> pad)."""
    H = sd["rnn.weight_hh_l0"].shape[1]
    E = sd["rnn.weight_ih_l0"].shape[1]
    cls = nn.GRU if _cell_of(sd) == "gru" else nn.LSTM
    rnn = cls(E, H, num_layers=2, bidirectional=True, batch_first=True).to(dtype)
    rnn.load_state_dict({k[4:]: v.to(dtype) for k, v in sd.items() if k.startswith("rnn.")})
    rnn.eval()
    packed = pack_padded_sequence(embed, lengths, batch_first=True)  # raises if unsorted
    out, _ = rnn(packed)
    out, _ = pad_packed_sequence(out, batch_first=True)
    return out


def _gru_direction(x, lengths, w_ih, w_hh, b_ih, b_hh, reverse):
    """One direction of one layer from the cell equations (gate rows r, z, n):
    r = s(gi_r+gh_r), z = s(gi_z+gh_z), n = tanh(gi_n + r*gh_n), h' = (1-z)n + z h.
    Padded steps produce zeros and do not advance the state (packed semantics)."""
    B, T, _ = x.shape
    H = w_hh.shape[1]
    out = x.new_zeros(B, T, H)
    for b in range(B):
        L = int(lengths[b])
        h = x.new_zeros(H)
        steps = range(L - 1, -1, -1) if reverse else range(L)
        for t in steps:
            gi = w_ih @ x[b, t] + b_ih
            gh = w_hh @ h + b_hh
            r = torch.sigmoid(gi[:H] + gh[:H])
            z = torch.sigmoid(gi[H:2 * H] + gh[H:2 * H])
            n = torch.tanh(gi[2 * H:] + r * gh[2 * H:])
            h = (1 - z) * n + z * h
            out[b, t] = h
    return out


def _lstm_direction(x, lengths, w_ih, w_hh, b_ih, b_hh, reverse):
    """nn.LSTM cell equations, gate rows i, f, g, o."""
    B, T, _ = x.shape
    H = w_hh.shape[1]
    out = x.new_zeros(B, T, H)
    for b in range(B):
        L = int(lengths[b])
        h = x.new_zeros(H)
        c = x.new_zeros(H)
        steps = range(L - 1, -1, -1) if reverse else range(L)
        for t in steps:
            g = w_ih @ x[b, t] + b_ih + w_hh @ h + b_hh
            i = torch.sigmoid(g[:H]); f = torch.sigmoid(g[H:2 * H])
            gg = torch.tanh(g[2 * H:3 * H]); o = torch.sigmoid(g[3 * H:])
            c = f * c + i * gg
            h = o * torch.tanh(c)
            out[b, t] = h
    return out


def _rnn_equations(sd, embed, lengths, dtype):
    fn = _gru_direction if _cell_of(sd) == "gru" else _lstm_direction
    x = embed
    for layer in range(2):
        outs = []
        for sfx, rev in (("", False), ("_reverse", True)):
            p = [sd[f"rnn.{n}_l{layer}{sfx}"].to(dtype)
                 for n in ("weight_ih", "weight_hh", "bias_ih", "bias_hh")]
            outs.append(fn(x, lengths, *p, reverse=rev))
        x = torch.cat(outs, dim=-1)
    return x


def predictor_forward(sd, i, x, dtype):
    """ArticulatorPredictor.forward, models.py:23-33 -> (..., 2, n_samples) pre-sigmoid."""
    p = f"predictors.{i}."
    g = lambda k: sd[p + k].to(dtype)
    h = F.layer_norm(x, (x.shape[-1],), g("linear.0.weight"), g("linear.0.bias"), LN_EPS)
    h = F.relu(F.linear(h, g("linear.1.weight"), g("linear.1.bias")))
    h = F.layer_norm(h, (256,), g("linear.3.weight"), g("linear.3.bias"), LN_EPS)
    h = F.relu(F.linear(h, g("linear.4.weight"), g("linear.4.bias")))
    h = F.layer_norm(h, (256,), g("linear.6.weight"), g("linear.6.bias"), LN_EPS)
    xs = F.linear(h, g("x_coords.weight"), g("x_coords.bias"))
    ys = F.linear(h, g("y_coords.weight"), g("y_coords.bias"))
    return torch.stack([xs, ys], dim=-2)


@torch.no_grad()
def artspeech_forward(sd, x, lengths, dtype=torch.float32, rnn="stock", return_hidden=False):
    """ArtSpeech.forward(x, lengths), models.py:126-145.

    x: (B, T) int64 tokens; lengths: sorted-descending ints (list or CPU tensor).
    Returns (B, max(lengths), N, 2, n_samples) in ``dtype``.
    """
    lengths = torch.as_tensor(lengths, dtype=torch.int64).cpu()
    embed = F.embedding(x.cpu(), sd["embedding.weight"].to(dtype))                     # :135
    if rnn == "stock":
        rnn_out = _rnn_stock(sd, embed, lengths, dtype)                                # :136-138
    else:
        if any(int(lengths[i]) < int(lengths[i + 1]) for i in range(len(lengths) - 1)):
            raise RuntimeError("`lengths` array must be sorted in decreasing order")
        rnn_out = _rnn_equations(sd, embed, lengths, dtype)[:, : int(lengths.max())]
    lin = F.relu(F.linear(rnn_out, sd["linear.0.weight"].to(dtype), sd["linear.0.bias"].to(dtype)))  # :140
    n_art = n_articulators_of(sd)
    out = torch.stack([predictor_forward(sd, i, lin, dtype) for i in range(n_art)], dim=2)  # :141-143
    out = torch.sigmoid(out)                                                           # :145
    if return_hidden:
        return out, rnn_out, lin
    return out


@torch.no_grad()
def simple_artspeech_forward(sd, x, dtype=torch.float32):
    """SimpleArtSpeech.forward(x, lengths), models.py:84-96 (eval mode: dropout is the identity)."""
    embed = F.embedding(x.cpu(), sd["embedding.weight"].to(dtype))                      # :91
    lin = F.relu(F.linear(embed, sd["linear.0.weight"].to(dtype), sd["linear.0.bias"].to(dtype)))  # :92
    n_art = n_articulators_of(sd)
    out = torch.stack([predictor_forward(sd, i, lin, dtype) for i in range(n_art)], dim=2)   # :93-95
    return torch.sigmoid(out)                                                           # :97


@torch.no_grad()
def pc_artspeech_forward(sd, x, lengths, dtype=torch.float32, rnn="stock"):
    """PrincipalComponentsArtSpeech.forward, principal_components/models/rnn.py:85-109 (predictor :11-33)."""
    lengths = torch.as_tensor(lengths, dtype=torch.int64).cpu()
    embed = F.embedding(x.cpu(), sd["embedding.weight"].to(dtype))                      # :94
    rnn_out = _rnn_stock(sd, embed, lengths, dtype) if rnn == "stock" else \
        _rnn_equations(sd, embed, lengths, dtype)[:, : int(lengths.max())]              # :95-104
    g = lambda k: sd[k].to(dtype)
    h = F.relu(F.linear(rnn_out, g("linear.0.weight"), g("linear.0.bias")))             # :106
    p = "predictor.linear."
    h = F.layer_norm(h, (h.shape[-1],), g(p + "0.weight"), g(p + "0.bias"), LN_EPS)     # :21
    h = F.relu(F.linear(h, g(p + "1.weight"), g(p + "1.bias")))                         # :22-23
    h = F.layer_norm(h, (h.shape[-1],), g(p + "3.weight"), g(p + "3.bias"), LN_EPS)     # :24
    h = F.relu(F.linear(h, g(p + "4.weight"), g(p + "4.bias")))                         # :25-26
    h = F.layer_norm(h, (h.shape[-1],), g(p + "6.weight"), g(p + "6.bias"), LN_EPS)     # :27
    return torch.tanh(F.linear(h, g(p + "7.weight"), g(p + "7.bias")))                  # :28, :107
