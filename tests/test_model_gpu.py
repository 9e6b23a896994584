"""GPU parity of the drop-in ArtSpeech against the reference's golden outputs and the oracle.
Gates (BASELINE.json north_star): contours within 2e-3 max-abs in fast (bf16/tf32) mode and
1e-5 in fp32 mode, against the reference fp32 forward."""
import os

import numpy as np
import pytest
import torch

from artspeech_b200 import synthetic

pytestmark = pytest.mark.gpu

TOL = {"fast": 2e-3, "fp32": 1e-5}


@pytest.fixture(scope="module")
def mods():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from artspeech_b200 import models
    from oracle import model as omodel
    return models, omodel


def _build(models, sd, precision, n_art=10, vocab=45):
    net = models.ArtSpeech(vocab, n_art, precision=precision)
    net.load_state_dict(sd)
    return net.cuda().eval()


@pytest.mark.parametrize("precision", ["fast", "fp32"])
@pytest.mark.parametrize("affine", [False, True])
def test_golden_small(mods, golden_dir, precision, affine):
    models, omodel = mods
    g = np.load(os.path.join(golden_dir, "model_small.npz"))
    sd = omodel.make_state_dict(45, 10, seed=0)
    if affine:
        sd = omodel.randomize_affine(sd, seed=1)
    net = _build(models, sd, precision)
    x = torch.from_numpy(g["x"]).cuda()
    with torch.no_grad():
        out = net(x, torch.from_numpy(g["lengths"]))
    want = g["out32_affine" if affine else "out32"]
    assert out.shape == want.shape and out.dtype == torch.float32 and out.is_contiguous()
    err = np.abs(out.cpu().numpy() - want).max()
    print(f"golden_small[{precision}, affine={affine}] max-abs {err:.3e}")
    assert err <= TOL[precision]


@pytest.mark.parametrize("precision", ["fast", "fp32"])
def test_variable_lengths_vs_oracle(mods, precision):
    models, omodel = mods
    sd = omodel.randomize_affine(omodel.make_state_dict(45, 10, seed=0), seed=3)
    net = _build(models, sd, precision)
    B, T = 37, 61
    x = synthetic.tokens(B, T, 45, seed=5)
    L = synthetic.sorted_lengths(B, 7, 53, seed=5)       # max(lengths) = 53 < T: output is cut to 53
    with torch.no_grad():
        out = net(x.cuda(), L)
    want = omodel.artspeech_forward(sd, x, L, torch.float32)
    assert out.shape == want.shape == (B, 53, 10, 2, 50)
    err = (out.cpu() - want).abs().max().item()
    print(f"variable_lengths[{precision}] max-abs {err:.3e}")
    assert err <= TOL[precision]
    # padded frames carry the constant decode(ReLU(b_linear)) frame (H3)
    b = B - 1
    pad = out[b, int(L[b]):]
    assert (pad - pad[0]).abs().max().item() == 0.0


def test_config2_shape_fast(mods):
    """BASELINE config 2: batch 256, length 100, checked against the CPU reference path."""
    models, omodel = mods
    sd = omodel.make_state_dict(45, 10, seed=0)
    net = _build(models, sd, "fast")
    x = synthetic.tokens(256, 100, 45, seed=0)
    L = [100] * 256
    with torch.no_grad():
        out = net(x.cuda(), L)
    want = omodel.artspeech_forward(sd, x[:32], L[:32], torch.float32)
    err = (out[:32].cpu() - want).abs().max().item()
    print(f"config2 fast max-abs {err:.3e}")
    assert err <= TOL["fast"]
    # batch rows are independent: a sub-batch gives the same contours
    with torch.no_grad():
        sub = net(x[:8].cuda(), L[:8])
    assert (sub - out[:8]).abs().max().item() <= 1e-6


def test_config2_shape_fp32(mods):
    """BASELINE config 2 (batch 256 x length 100) in fp32 precision: the 1e-5 gate against the CPU
    reference path (rows are independent, the oracle runs 32 of them), full and ragged lengths."""
    models, omodel = mods
    sd = omodel.randomize_affine(omodel.make_state_dict(45, 10, seed=0), seed=7)
    net = _build(models, sd, "fp32")
    x = synthetic.tokens(256, 100, 45, seed=0)
    for L in ([100] * 256, synthetic.sorted_lengths(256, 50, 100, seed=3)):
        L = list(L)
        L[0] = 100
        with torch.no_grad():
            out = net(x.cuda(), L)
        rows = list(range(16)) + list(range(240, 256))
        for lo, hi in ((0, 16), (240, 256)):
            want = omodel.artspeech_forward(sd, x[lo:hi], L[lo:hi], torch.float32)
            err = (out[lo:hi, :want.shape[1]].cpu() - want).abs().max().item()
            print(f"config2 fp32 rows {lo}:{hi} max-abs {err:.3e}")
            assert err <= TOL["fp32"]


@pytest.mark.parametrize("gamma,beta", [(0.25, 0.1), (1.0, 0.3), (2.0, 0.5), (4.0, 1.0)])
def test_fast_mode_margin_over_layernorm_scales(mods, gamma, beta):
    """Trained-like weight scales: LayerNorm gains multiplied by `gamma`, biases perturbed by N(0, beta^2)
    (round-1 VERDICT: nothing explored where fast precision breaks).  The fast-precision error grows linearly
    with the gain of the last LayerNorm (it multiplies the head logits); measured on B200 with every GEMM operand
    in fp16: 1.6e-4 / 6.3e-4 / 1.3e-3 / 2.4e-3 at gamma = 0.25 / 1 / 2 / 4 (2.7e-4 / 1.1e-3 / 2.3e-3 / 4.5e-3 with
    the bf16 recurrent operands of round 1), i.e. the 2e-3 gate of the north star (stated for random-init weights,
    gamma = 1) holds up to gamma ~ 3; fp32 precision (split operands) gives 7e-7 ... 8e-6 and holds 1e-5 over the
    whole range.  The test pins those margins: a checkpoint with larger gains needs fp32 precision."""
    models, omodel = mods
    sd = omodel.make_state_dict(45, 10, seed=0)
    g = torch.Generator().manual_seed(11)
    for k, v in sd.items():
        part = k.split(".")
        if k.startswith("predictors.") and part[2] == "linear" and part[3] in ("0", "3", "6"):
            if k.endswith("weight"):
                sd[k] = v * gamma * (1 + 0.2 * torch.randn(v.shape, generator=g))
            else:
                sd[k] = v + beta * torch.randn(v.shape, generator=g)
    x = synthetic.tokens(16, 40, 45, seed=2)
    L = synthetic.sorted_lengths(16, 10, 40, seed=2)
    want = omodel.artspeech_forward(sd, x, L, torch.float32)
    want64 = omodel.artspeech_forward(sd, x, L, torch.float64)
    errs = {}
    for precision in ("fast", "fp32"):
        net = _build(models, sd, precision)
        with torch.no_grad():
            out = net(x.cuda(), L).cpu()
        errs[precision] = (out - want).abs().max().item()
        errs[precision + "_vs64"] = (out.double() - want64).abs().max().item()
    saturated = float(((want < 1e-3) | (want > 1 - 1e-3)).float().mean())
    print(f"margin[gamma={gamma}, beta={beta}] fast {errs['fast']:.3e} fp32 {errs['fp32']:.3e} "
          f"(vs fp64: {errs['fast_vs64']:.3e} / {errs['fp32_vs64']:.3e}); saturated outputs {saturated:.3f}")
    assert errs["fp32"] <= TOL["fp32"]
    assert errs["fast"] <= (TOL["fast"] if gamma <= 2.0 else 1e-3 * gamma)


def test_out_of_range_tokens_raise_index_error(mods):
    """nn.Embedding raises IndexError for ids outside [0, vocab) (models.py:135).  Host tensors are
    checked before launch; device tensors are flagged by the kernels (tcgen05 recurrent kernel and the
    gather path) and the error surfaces at check_tokens() / the next call - never a silent clamp."""
    models, omodel = mods
    sd = omodel.make_state_dict(45, 10, seed=0)
    x = synthetic.tokens(4, 9, 45, seed=1)
    bad = x.clone()
    bad[2, 3] = 45
    neg = x.clone()
    neg[0, 0] = -1
    for precision in ("fast", "fp32"):        # fast: tokens read by the recurrent kernel; fp32: gather kernel
        net = _build(models, sd, precision)
        with torch.no_grad():
            with pytest.raises(IndexError):
                net(bad, [9, 9, 9, 9])                     # host tensor: raised before anything runs
            good = net(x.cuda(), [9, 9, 9, 9])
            net.check_tokens()                             # nothing to report
            net(bad.cuda(), [9, 9, 9, 9])                  # device tensor: recorded by the kernel ...
            with pytest.raises(IndexError, match="45"):
                net.check_tokens()                         # ... and raised here
            net.check_tokens()                             # reported once
            net(neg.cuda(), [9, 9, 9, 9])
            torch.cuda.synchronize()
            with pytest.raises(IndexError):
                net(x.cuda(), [9, 9, 9, 9])                # lazily: the next call raises
            again = net(x.cuda(), [9, 9, 9, 9])
        assert torch.equal(good, again)


def test_standalone_articulator_predictor(mods, golden_dir):
    """ArticulatorPredictor.forward on its own (models.py:23-33; the transformer heads,
    transformer/models.py:337-339) against the reference class's own outputs (predictor.npz)."""
    models, omodel = mods
    g = np.load(os.path.join(golden_dir, "predictor.npz"))
    for tag, (c, d) in {"a": (64, 50), "b": (256, 30), "c": (128, 50)}.items():
        for precision in ("fast", "fp32"):
            torch.manual_seed(0)
            net = models.ArticulatorPredictor(c, d, precision=precision).eval()
            gen = torch.Generator().manual_seed(17)
            with torch.no_grad():
                for k, v in net.state_dict().items():
                    if k.startswith("linear.") and v.dim() == 1 and k.split(".")[1] in ("0", "3", "6"):
                        v.add_(0.3 * torch.randn(v.shape, generator=gen))
            sd = net.state_dict()
            keys = sorted(sd.keys())
            assert keys == list(g[f"keys_{tag}"])
            sums = np.array([sd[k].double().sum().item() for k in keys])
            assert np.allclose(sums, g[f"sums_{tag}"], rtol=0, atol=1e-9)
            net = net.cuda()
            x = torch.from_numpy(g[f"x_{tag}"]).cuda()
            with torch.no_grad():
                out = net(x)
            want = g[f"out32_{tag}"]
            assert tuple(out.shape) == want.shape, (tag, out.shape, want.shape)
            scale = max(1.0, float(np.abs(want).max()))
            err = float(np.abs(out.cpu().numpy() - want).max()) / scale
            print(f"predictor[{tag}, {precision}] max-abs/scale {err:.3e} (|out| max {scale:.2f})")
            assert err <= (2e-3 if precision == "fast" else 1e-5)
    # under autograd: same values, gradients of the torch graph of the same parameters
    torch.manual_seed(0)
    net = models.ArticulatorPredictor(64, 50, precision="fp32").cuda()
    x = torch.randn(3, 7, 64, device="cuda", requires_grad=True)
    out = net(x)
    out.square().mean().backward()
    ref = net._graph(x.detach())
    assert (out.detach() - ref).abs().max().item() <= 1e-5 * max(1.0, ref.abs().max().item())
    assert x.grad is not None and all(p.grad is not None for p in net.parameters())
    with pytest.raises(RuntimeError):
        net(torch.randn(3, 7, 32, device="cuda"))


def test_two_devices_in_one_process(mods):
    """cudaFuncSetAttribute opt-ins are per device: the model and the metrics on cuda:1 after cuda:0."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    models, omodel = mods
    from artspeech_b200.metrics import p2cp_distance
    sd = omodel.make_state_dict(45, 10, seed=0)
    x = synthetic.tokens(8, 20, 45, seed=1)
    tgt = synthetic.contours(8 * 20, 10, 50, seed=2).view(8, 20, 10, 2, 50)
    outs = []
    for d in (0, 1):
        net = models.ArtSpeech(45, 10)
        net.load_state_dict(sd)
        net = net.to(f"cuda:{d}").eval()
        with torch.no_grad():
            o = net(x.to(f"cuda:{d}"), [20] * 8)
        p = p2cp_distance(o, tgt.to(f"cuda:{d}"))
        outs.append((o.cpu(), p.cpu()))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])


def test_decoder_implementations_agree(mods):
    """The three fused decoders of fast precision - tcgen05 with fp16 operands (default), tcgen05 with tf32
    operands (round 1) and mma.sync tf32 - against each other and the fp32 oracle, on a batch that leaves a
    partly filled 128-frame tile and with perturbed LayerNorm affines."""
    models, omodel = mods
    sd = omodel.randomize_affine(omodel.make_state_dict(45, 10, seed=0), seed=9)
    net = _build(models, sd, "fast")
    B, T = 41, 37
    x = synthetic.tokens(B, T, 45, seed=3)
    L = synthetic.sorted_lengths(B, 5, T, seed=3)
    want = omodel.artspeech_forward(sd, x[:8], L[:8], torch.float32)
    outs = {}
    for impl in ("tcgen05", "tcgen05-tf32", "mma"):
        net.set_decoder(impl)
        with torch.no_grad():
            outs[impl] = net(x.cuda(), L).cpu()
        err = (outs[impl][:8, :want.shape[1]] - want).abs().max().item()
        print(f"decoder[{impl}] max-abs vs fp32 oracle {err:.3e}")
        assert err <= TOL["fast"] and not torch.isnan(outs[impl]).any()
    assert (outs["tcgen05"] - outs["tcgen05-tf32"]).abs().max().item() <= 1e-3
    assert (outs["tcgen05"] - outs["mma"]).abs().max().item() <= 1e-3
    net.set_decoder("tcgen05")


def test_fp32_decoder_implementations_agree(mods):
    """fp32 precision: the tcgen05 decoder with SPLIT fp16 operands (hi + lo, three MMAs per product; default)
    against the fp32 FFMA decoder and the fp32 oracle, partly filled tile, perturbed LayerNorm affines."""
    models, omodel = mods
    sd = omodel.randomize_affine(omodel.make_state_dict(45, 10, seed=0), seed=9)
    net = _build(models, sd, "fp32")
    B, T = 41, 37
    x = synthetic.tokens(B, T, 45, seed=3)
    L = synthetic.sorted_lengths(B, 5, T, seed=3)
    want = omodel.artspeech_forward(sd, x[:8], L[:8], torch.float32)
    outs = {}
    for impl in ("tcgen05", "mma"):
        net.set_decoder(impl)
        with torch.no_grad():
            outs[impl] = net(x.cuda(), L).cpu()
        err = (outs[impl][:8, :want.shape[1]] - want).abs().max().item()
        print(f"fp32 decoder[{impl}] max-abs vs fp32 oracle {err:.3e}")
        assert err <= TOL["fp32"] and not torch.isnan(outs[impl]).any()
    assert (outs["tcgen05"] - outs["mma"]).abs().max().item() <= 5e-6
    with pytest.raises(ValueError):
        net.set_decoder("tcgen05-tf32")      # the tf32 kernel exists in fast precision only
    net.set_decoder("tcgen05")


def test_fp32_recurrent_implementations_agree(mods):
    """fp32 precision: the tcgen05 recurrent kernel with split fp16 operands (W_hh and h as hi + lo, three MMAs
    per product; default) against the FFMA kernel and the fp32 oracle, ragged lengths, 300 steps."""
    models, omodel = mods
    sd = omodel.make_state_dict(45, 3, seed=4)
    net = _build(models, sd, "fp32", n_art=3)
    B, T = 13, 300
    x = synthetic.tokens(B, T, 45, seed=6)
    L = synthetic.sorted_lengths(B, 7, T, seed=6)
    want = omodel.artspeech_forward(sd, x[:4], L[:4], torch.float32)
    outs = {}
    for impl in ("tcgen05", "mma"):
        net.set_recurrent(impl)
        with torch.no_grad():
            outs[impl] = net(x.cuda(), L).cpu()
        err = (outs[impl][:4, :want.shape[1]] - want).abs().max().item()
        print(f"fp32 recurrent[{impl}] max-abs vs fp32 oracle {err:.3e}")
        assert err <= TOL["fp32"] and not torch.isnan(outs[impl]).any()
    assert (outs["tcgen05"] - outs["mma"]).abs().max().item() <= 5e-6
    net.set_recurrent("tcgen05")


@pytest.mark.parametrize("precision", ["fast", "fp32"])
@pytest.mark.parametrize("n_samples", [40, 64])
def test_other_contour_sizes(mods, precision, n_samples):
    """ArtSpeech(n_samples != 50) (models.py:100-101): the tcgen05 decoder and projection are specialised for 50
    points, other sizes run the mma.sync decoder / projection; the recurrent layers are the same kernels."""
    models, omodel = mods
    sd = omodel.make_state_dict(45, 4, n_samples=n_samples, seed=5)
    net = models.ArtSpeech(45, 4, n_samples=n_samples, precision=precision)
    net.load_state_dict(sd)
    net = net.cuda().eval()
    B, T = 9, 31
    x = synthetic.tokens(B, T, 45, seed=8)
    L = synthetic.sorted_lengths(B, 3, T, seed=8)
    with torch.no_grad():
        out = net(x.cuda(), L).cpu()
    want = omodel.artspeech_forward(sd, x, L, torch.float32)
    assert out.shape == want.shape == (B, max(L), 4, 2, n_samples)
    err = (out - want).abs().max().item()
    print(f"n_samples={n_samples}[{precision}] max-abs {err:.3e}")
    assert err <= TOL[precision]
    with pytest.raises(ValueError):
        net.set_decoder("tcgen05")


def test_errors_and_state_dict(mods):
    models, omodel = mods
    sd = omodel.make_state_dict(45, 2, seed=0)
    net = models.ArtSpeech(45, 2)
    assert sorted(net.state_dict().keys()) == sorted(sd.keys())
    for k, v in net.state_dict().items():
        assert v.shape == sd[k].shape
    net.load_state_dict(sd)
    assert net.total_parameters == sum(v.numel() for v in sd.values())
    net = net.cuda().eval()
    x = synthetic.tokens(3, 9, 45).cuda()
    with torch.no_grad():
        with pytest.raises(RuntimeError):
            net(x, [5, 9, 3])              # unsorted, like pack_padded_sequence
        with pytest.raises(RuntimeError):
            net(x, [9, 5, 0])              # zero length
        a = net(x, [9, 5, 3])
        # weights changed in place -> the native copy is rebuilt
        net.linear[0].bias.add_(1.0)
        b = net(x, [9, 5, 3])
    assert (a - b).abs().max().item() > 1e-4
    out = net(x, [9, 5, 3])                # autograd enabled: the training path (see test_training_step_*)
    assert out.requires_grad and (out.detach() - b).abs().max().item() <= 2e-3   # fp32 training path vs fast inference


def test_tcgen05_recurrent_kernel_matches_register_resident_kernel(mods):
    """gru_umma.cu (W_hh in tensor memory, tcgen05.mma, 4/8/16 rows per CTA) against gru.cu."""
    import torch
    from artspeech_b200 import models, synthetic
    from oracle import model as omodel
    sd = omodel.randomize_affine(omodel.make_state_dict(45, 10, seed=0), seed=2)
    net = models.ArtSpeech(45, 10, precision="fast")
    net.load_state_dict(sd)
    net = net.cuda().eval()
    for (B, T) in [(5, 23), (256, 40), (700, 12), (1300, 9)]:     # 4, 4, 8 and 16 rows per CTA
        x = synthetic.tokens(B, T, 45, seed=B)
        L = synthetic.sorted_lengths(B, max(1, T // 4), T, seed=B)
        with torch.no_grad():
            net.set_recurrent("mma")
            a = net(x.cuda(), L)
            net.set_recurrent("tcgen05")       # the fast-mode default
            b = net(x.cuda(), L)
        torch.cuda.synchronize()
        assert not torch.isnan(b).any()
        # the register-resident kernel keeps bf16 operands (8 significand bits), the tcgen05 kernel fp16 (11): the two
        # differ by the former's rounding, and the tcgen05 kernel is the one closer to the fp32 oracle
        assert float((a - b).abs().max()) <= 2e-3, (B, T, float((a - b).abs().max()))
        nb = min(B, 6)
        want = omodel.artspeech_forward(sd, x[:nb], L[:nb], torch.float32)
        err_b = float((b[:nb, :want.shape[1]].cpu() - want).abs().max())
        err_a = float((a[:nb, :want.shape[1]].cpu() - want).abs().max())
        print(f"recurrent kernels [{B}x{T}] vs fp32 oracle: tcgen05 (fp16) {err_b:.3e}, mma.sync (bf16) {err_a:.3e}")
        assert err_b <= 1.2e-3 and err_a <= 2e-3


@pytest.mark.parametrize("precision,tol", [("fast", 2e-3), ("fp32", 1e-5)])
def test_config4_long_sequence(mods, precision, tol):
    """BASELINE configs[3]: batch 64, length 1024 - 1024 dependent recurrent steps per direction.
    Rows are independent, so the oracle (fp32 CPU forward) is run on 6 of the 64 rows."""
    import torch
    from artspeech_b200 import models, synthetic
    from oracle import model as omodel
    sd = omodel.randomize_affine(omodel.make_state_dict(45, 10, seed=0), seed=4)
    net = models.ArtSpeech(45, 10, precision=precision)
    net.load_state_dict(sd)
    net = net.cuda().eval()
    B, T = 64, 1024
    x = synthetic.tokens(B, T, 45, seed=44)
    L = synthetic.sorted_lengths(B, 300, T, seed=44)
    with torch.no_grad():
        out = net(x.cuda(), L)
    torch.cuda.synchronize()
    assert tuple(out.shape) == (B, T, 10, 2, 50) and not torch.isnan(out).any()
    rows = [0, 1, 17, 40, 62, 63]
    for b in rows:
        lb = int(L[b])
        want = omodel.artspeech_forward(sd, x[b:b + 1], [lb], torch.float32)
        err = float((out[b:b + 1, :lb].cpu() - want).abs().max())
        assert err <= tol, (precision, b, err)
    # padded steps hold the constant padded frame (the decoder applied to a zero GRU output)
    b = 63
    lb = int(L[b])
    if lb < T:
        pad = out[b, lb:]
        assert float((pad - pad[0:1]).abs().max()) == 0.0


def test_lstm_cell_matches_nn_lstm_oracle(mods):
    """RNNType.LSTM (phoneme_to_articulation/__init__.py:47-49): the 4-gate instantiation of the
    tcgen05 recurrent kernel against the nn.LSTM-based oracle (stock cell and cell equations)."""
    import torch
    from artspeech_b200 import models, synthetic
    from oracle import model as omodel
    sd = omodel.randomize_affine(omodel.make_state_dict(45, 10, seed=0, cell="lstm"), seed=5)
    net = models.ArtSpeech(45, 10, precision="fast", rnn="lstm")
    assert sorted(net.state_dict()) == sorted(sd)
    net.load_state_dict(sd)
    net = net.cuda().eval()
    for (B, T) in [(1, 1), (5, 23), (256, 40), (700, 12), (1300, 9)]:   # 4, 4, 4, 8 and 16 rows per CTA
        x = synthetic.tokens(B, T, 45, seed=B)
        L = synthetic.sorted_lengths(B, max(1, T // 4), T, seed=B)
        with torch.no_grad():
            out = net(x.cuda(), L)
        torch.cuda.synchronize()
        assert not torch.isnan(out).any()
        nb = min(B, 6)
        want = omodel.artspeech_forward(sd, x[:nb], L[:nb], torch.float32)
        err = float((out[:nb, :want.shape[1]].cpu() - want).abs().max())
        assert err <= 2e-3, (B, T, err)
        if B == 5:
            eq = omodel.artspeech_forward(sd, x[:nb], L[:nb], torch.float64, rnn="equations")
            assert float((out[:nb, :eq.shape[1]].cpu().double() - eq).abs().max()) <= 2e-3
    # the mma.sync decoder / projection pair works with the 4-gate layout as well
    with torch.no_grad():
        a = net(x.cuda(), L)
        net.set_decoder("mma")
        b = net(x.cuda(), L)
        net.set_decoder("tcgen05")
    assert float((a - b).abs().max()) <= 1e-3
    with pytest.raises(ValueError):
        net.set_recurrent("mma")           # GRU-only kernel


def test_lstm_fp32_precision(mods):
    """RNNType.LSTM in fp32 precision (lstm_fp32_kernel: W_hh^T in shared memory + registers) meets the
    1e-5 gate against the nn.LSTM-based oracle, variable lengths, partly filled CTAs."""
    models, omodel = mods
    sd = omodel.randomize_affine(omodel.make_state_dict(45, 10, seed=0, cell="lstm"), seed=5)
    net = models.ArtSpeech(45, 10, precision="fp32", rnn="lstm")
    net.load_state_dict(sd)
    net = net.cuda().eval()
    for (B, T) in [(1, 1), (5, 23), (37, 61), (64, 200)]:
        x = synthetic.tokens(B, T, 45, seed=B)
        L = synthetic.sorted_lengths(B, max(1, T // 4), T, seed=B)
        L[0] = T
        with torch.no_grad():
            out = net(x.cuda(), L)
        nb = min(B, 5)
        want = omodel.artspeech_forward(sd, x[:nb], L[:nb], torch.float32)
        err = float((out[:nb, :want.shape[1]].cpu() - want).abs().max())
        print(f"lstm fp32 [{B}x{T}] max-abs {err:.3e}")
        assert err <= TOL["fp32"], (B, T, err)
        assert not torch.isnan(out).any()


def test_lstm_config4_long_sequence(mods):
    """configs[3] with the LSTM cell: 1024 dependent steps keep the bf16-operand recurrence in gate."""
    import torch
    from artspeech_b200 import models, synthetic
    from oracle import model as omodel
    sd = omodel.randomize_affine(omodel.make_state_dict(45, 10, seed=0, cell="lstm"), seed=6)
    net = models.ArtSpeech(45, 10, precision="fast", rnn="lstm")
    net.load_state_dict(sd)
    net = net.cuda().eval()
    B, T = 64, 1024
    x = synthetic.tokens(B, T, 45, seed=45)
    L = synthetic.sorted_lengths(B, 300, T, seed=45)
    with torch.no_grad():
        out = net(x.cuda(), L)
    torch.cuda.synchronize()
    assert tuple(out.shape) == (B, T, 10, 2, 50) and not torch.isnan(out).any()
    for b in [0, 31, 63]:
        lb = int(L[b])
        want = omodel.artspeech_forward(sd, x[b:b + 1], [lb], torch.float32)
        err = float((out[b:b + 1, :lb].cpu() - want).abs().max())
        assert err <= 2e-3, (b, err)


@pytest.mark.parametrize("precision", ["fast", "fp32"])
def test_simple_artspeech_golden(mods, golden_dir, precision):
    """SimpleArtSpeech (models.py:53-96) as a lookup in the per-token contour table, against the
    reference's own output (tests/golden/model_small.npz: out_simple) and the fp64 oracle."""
    models, omodel = mods
    g = np.load(os.path.join(golden_dir, "model_small.npz"))
    torch.manual_seed(0)
    net = models.SimpleArtSpeech(45, 10, precision=precision)
    sd = {k: v.detach().clone() for k, v in net.state_dict().items()}
    net = net.cuda().eval()
    x = torch.from_numpy(g["x"]).cuda()
    with torch.no_grad():
        out = net(x, torch.from_numpy(g["lengths"]))
    assert out.shape == g["out_simple"].shape and out.dtype == torch.float32 and out.is_contiguous()
    err = np.abs(out.cpu().numpy() - g["out_simple"]).max()
    print(f"simple_artspeech[{precision}] max-abs vs reference {err:.3e}")
    assert err <= TOL[precision]
    # perturbed LayerNorm affines + a full-vocabulary batch; every row equals the row of its token
    sd2 = omodel.randomize_affine(sd, seed=4)
    net.load_state_dict(sd2)
    x2 = synthetic.tokens(7, 33, 45, seed=9)
    with torch.no_grad():
        out2 = net(x2.cuda(), [33] * 7).cpu()
    want = omodel.simple_artspeech_forward(sd2, x2, torch.float64)
    assert (out2.double() - want).abs().max().item() <= TOL[precision]
    flat, tok = out2.view(-1, 1000), x2.view(-1)
    first = {}
    for r, t in enumerate(tok.tolist()):
        first.setdefault(t, r)
        assert torch.equal(flat[r], flat[first[t]])
    # out-of-range token: nn.Embedding raises IndexError (models.py:63); the kernel writes a NaN row and
    # records the id, check_tokens() raises; a host tensor is refused before launch
    bad = x2.clone(); bad[0, 0] = 45
    with torch.no_grad():
        ob = net(bad.cuda(), None).cpu()
        with pytest.raises(IndexError, match="45"):
            net.check_tokens()
        with pytest.raises(IndexError):
            net(bad, None)
    assert torch.isnan(ob[0, 0]).all() and not torch.isnan(ob[0, 1:]).any()


def test_gather_rows_abi(mods):
    from artspeech_b200 import _lib
    table = torch.randn(45, 256, device="cuda")
    tok = torch.randint(0, 45, (1000,), device="cuda")
    out = torch.empty(1000, 256, device="cuda")
    _lib.check(_lib.lib().asb_gather_rows(_lib.ptr(table), _lib.ptr(tok), 1000, 45, 256, _lib.ptr(out),
                                          _lib.stream_ptr(out.device)))
    assert torch.equal(out, table[tok])
    with pytest.raises(ValueError):
        _lib.check(_lib.lib().asb_gather_rows(_lib.ptr(table), _lib.ptr(tok), 1000, 45, 255, _lib.ptr(out),
                                              _lib.stream_ptr(out.device)))


@pytest.mark.parametrize("B,T,n_art,vocab", [(1, 1, 10, 45), (17, 5, 3, 45), (5, 40, 2, 3000), (130, 3, 1, 7)])
def test_edge_shapes_vs_oracle(mods, B, T, n_art, vocab):
    """Single frame, batch sizes that leave partly filled CTAs / tiles, articulator counts other than
    10, and a vocabulary whose layer-0 table does not fit beside the recurrent kernel's shared memory
    (the gather path).  Buffers are reused across calls of different shape: padded steps must not leak."""
    models, omodel = mods
    sd = omodel.randomize_affine(omodel.make_state_dict(vocab, n_art, seed=2), seed=5)
    net = _build(models, sd, "fast", n_art=n_art, vocab=vocab)
    with torch.no_grad():
        net(synthetic.tokens(B + 3, T + 9, vocab, seed=1).cuda(), [T + 9] * (B + 3))   # dirties the workspace
    x = synthetic.tokens(B, T, vocab, seed=8)
    L = synthetic.sorted_lengths(B, 1, T, seed=8) if T > 1 else [1] * B
    L[0] = T
    with torch.no_grad():
        out = net(x.cuda(), L)
    want = omodel.artspeech_forward(sd, x, L, torch.float32)
    assert out.shape == want.shape == (B, T, n_art, 2, 50)
    err = (out.cpu() - want).abs().max().item()
    print(f"edge[{B}x{T}, {n_art} articulators, vocab {vocab}] max-abs {err:.3e}")
    assert err <= TOL["fast"]


def test_training_step_gradients_match_reference(mods, golden_dir):
    """train_phoneme_to_articulation.py:84-96 on the drop-in module: forward with saved activations and
    HAND-WRITTEN backward kernels (artspeech_b200/training.py; cuDNN disabled to prove it), gradients equal
    to the reference model's own autograd gradients (tests/golden/model_grad.npz), and an optimizer step
    changes what the kernels compute next."""
    models, omodel = mods
    g = np.load(os.path.join(golden_dir, "model_grad.npz"))
    torch.manual_seed(0)
    net = models.ArtSpeech(45, 10, precision="fp32").cuda().train()
    x = torch.from_numpy(g["x"]).cuda()
    lengths = torch.from_numpy(g["lengths"])
    w = torch.from_numpy(g["w"]).cuda()
    opt = torch.optim.SGD(net.parameters(), lr=0.05)
    opt.zero_grad()
    from artspeech_b200 import _lib
    with torch.backends.cudnn.flags(enabled=False):
        l0 = _lib.launch_count()
        out = net(x, lengths)
        assert out.requires_grad
        (out * w).sum().backward()
        assert _lib.launch_count() - l0 > 200          # the step ran on this library's kernels
    with torch.no_grad():
        ref_out = net._native_forward(x, lengths)
    assert (out.detach() - ref_out).abs().max().item() <= 1e-5     # training forward == inference forward
    named = dict(net.named_parameters())
    keys = list(g["keys"])
    sums = np.array([named[k].grad.double().sum().item() for k in keys])
    abssums = np.array([named[k].grad.double().abs().sum().item() for k in keys])
    assert np.allclose(abssums, g["abssums64"], rtol=2e-4, atol=1e-6)
    assert np.abs(sums - g["sums64"]).max() <= 1e-4 * np.abs(g["abssums64"]).max()
    for k in ("linear.0.bias", "rnn.bias_hh_l0", "rnn.bias_ih_l1_reverse", "predictors.3.x_coords.bias"):
        want = g["grad64:" + k]
        err = np.abs(named[k].grad.cpu().numpy() - want).max()
        assert err <= 1e-4 * max(1.0, np.abs(want).max()), (k, err)
    before = out.detach().clone()
    opt.step()
    with torch.no_grad():
        after = net(x, lengths)                              # weights changed in place -> re-packed
    assert (after - before).abs().max().item() > 1e-5
    # fast precision trains too (the training path is fp32 whatever the inference precision is)
    net2 = models.ArtSpeech(45, 10, precision="fast").cuda().train()
    o2 = net2(x, lengths)
    o2.mean().backward()
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in net2.parameters())


def test_training_gradients_vs_fp64_oracle_ragged_batch(mods):
    """Every parameter's gradient against fp64 autograd through the oracle graph, on a batch that leaves
    partly filled CTAs / tiles and has ragged lengths (BPTT must respect each row's own length)."""
    models, omodel = mods
    torch.manual_seed(3)
    net = models.ArtSpeech(45, 3, precision="fp32").cuda().train()
    import copy
    twin = copy.deepcopy(net).cpu().double()        # the same module graph in fp64 torch ops: the gradient truth
    B, T = 11, 17
    x = synthetic.tokens(B, T, 45, seed=4)
    L = [17, 17, 16, 12, 12, 9, 5, 3, 2, 1, 1]
    w = torch.randn(B, T, 3, 2, 50, generator=torch.Generator().manual_seed(5))
    out = net(x.cuda(), L)
    (out * w.cuda()).sum().backward()
    want = twin._library_graph(x, L)
    (want * w.double()).sum().backward()
    assert (out.detach().cpu().double() - want.detach()).abs().max().item() <= 1e-5
    truth = dict(twin.named_parameters())
    for k, p in net.named_parameters():
        got, gw = p.grad.cpu().double(), truth[k].grad
        err = (got - gw).abs().max().item()
        assert err <= 2e-4 * max(1.0, gw.abs().max().item()), (k, err, gw.abs().max().item())


def test_training_lstm_cell_gradients_vs_fp64(mods):
    """The LSTM instantiation trains natively too (lstm_train_{fwd,bwd}_kernel): every gradient against fp64
    autograd of the same module graph, ragged batch, cuDNN disabled."""
    import copy
    models, omodel = mods
    torch.manual_seed(5)
    net = models.ArtSpeech(45, 2, precision="fp32", rnn="lstm").cuda().train()
    twin = copy.deepcopy(net).cpu().double()
    B, T = 9, 14
    x = synthetic.tokens(B, T, 45, seed=6)
    L = [14, 13, 13, 10, 7, 4, 2, 1, 1]
    w = torch.randn(B, T, 2, 2, 50, generator=torch.Generator().manual_seed(7))
    with torch.backends.cudnn.flags(enabled=False):
        out = net(x.cuda(), L)
        (out * w.cuda()).sum().backward()
    want = twin._library_graph(x, L)
    (want * w.double()).sum().backward()
    assert (out.detach().cpu().double() - want.detach()).abs().max().item() <= 1e-5
    truth = dict(twin.named_parameters())
    for k, p in net.named_parameters():
        got, gw = p.grad.cpu().double(), truth[k].grad
        err = (got - gw).abs().max().item()
        assert err <= 2e-4 * max(1.0, gw.abs().max().item()), (k, err, gw.abs().max().item())


def test_training_dropout(mods):
    """nn.GRU(dropout=p) between the layers (models.py:111) in train mode: a mask is applied (train != eval),
    torch.manual_seed reproduces it, eval mode ignores it, and the analytic gradient agrees with a central
    finite difference taken under the same mask."""
    models, omodel = mods
    torch.manual_seed(0)
    net = models.ArtSpeech(45, 2, dropout=0.5, precision="fp32").cuda()
    x = synthetic.tokens(6, 11, 45, seed=1).cuda()
    L = [11, 10, 7, 7, 4, 2]
    w = torch.randn(6, 11, 2, 2, 50, generator=torch.Generator().manual_seed(2)).cuda()
    net.eval()
    with torch.no_grad():
        e1 = net(x, L)
    net.train()
    torch.manual_seed(7); a = net(x, L)
    torch.manual_seed(7); b = net(x, L)
    torch.manual_seed(8); c = net(x, L)
    assert torch.equal(a.detach(), b.detach()) and not torch.equal(a.detach(), c.detach())
    assert (a.detach() - e1).abs().max().item() > 1e-4
    with torch.no_grad():
        torch.manual_seed(7)
        d = net(x, L)                                       # train mode without autograd still drops
    assert torch.equal(d, a.detach())
    (a * w).sum().backward()
    p = net.rnn.bias_ih_l0
    grad = p.grad.clone()
    k = int(grad.abs().argmax())
    eps = 1e-2
    vals = []
    for sgn in (+1, -1):
        with torch.no_grad():
            p[k] += sgn * eps
            torch.manual_seed(7)
            vals.append(float((net(x, L).double() * w.double()).sum()))
            p[k] -= sgn * eps
    fd = (vals[0] - vals[1]) / (2 * eps)
    assert abs(fd - float(grad[k])) <= 3e-2 * abs(fd) + 1e-3, (fd, float(grad[k]))


def test_simple_artspeech_training_step(mods):
    """SimpleArtSpeech under autograd: native forward values, gradients equal to those of the fp64
    oracle graph (models.py:84-96), and the per-token table follows an optimizer step."""
    models, omodel = mods
    torch.manual_seed(0)
    net = models.SimpleArtSpeech(45, 10, precision="fp32").cuda().train()
    sd = {k: v.detach().cpu().double().requires_grad_(True) for k, v in net.state_dict().items()}
    x = synthetic.tokens(4, 9, 45, seed=6)
    w = torch.randn(4, 9, 10, 2, 50, generator=torch.Generator().manual_seed(2))
    out = net(x.cuda(), [9] * 4)
    assert out.requires_grad
    (out * w.cuda()).sum().backward()
    want = omodel.simple_artspeech_forward.__wrapped__(sd, x, torch.float64)
    (want * w.double()).sum().backward()
    assert (out.detach().cpu().double() - want.detach()).abs().max().item() <= TOL["fp32"]
    for k, p in net.named_parameters():
        g, gw = p.grad.cpu().double(), sd[k].grad
        assert (g - gw).abs().max().item() <= 1e-4 * max(1.0, gw.abs().max().item()), k
    opt = torch.optim.SGD(net.parameters(), lr=0.1)
    opt.step()
    with torch.no_grad():
        after = net(x.cuda(), None)
    assert (after - out.detach()).abs().max().item() > 1e-5
