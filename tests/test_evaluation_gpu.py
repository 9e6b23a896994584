"""Device-resident evaluation step: fused accumulation kernel, CUDA-graph replay, TV with an
external upper incisor - each against the plain torch / oracle computation."""
import pytest
import torch

from artspeech_b200 import synthetic

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def setup():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from artspeech_b200 import evaluation, models
    torch.manual_seed(0)
    net = models.ArtSpeech(45, 10).cuda().eval()
    B, T = 19, 33
    x = synthetic.tokens(B, T, 45, seed=2).cuda()
    L = synthetic.sorted_lengths(B, 5, T, seed=2)
    tgt = synthetic.contours(B * T, 10, 50, seed=3).view(B, T, 10, 2, 50).cuda()
    return evaluation, net, x, L, tgt


def test_accumulate_kernel_matches_torch(setup):
    evaluation, net, x, L, tgt = setup
    res = evaluation.evaluate_batch(net, x, L, tgt)
    want = evaluation.accumulate(res["p2cp"], res["tv_pred"], res["tv_target"], L)
    assert torch.allclose(res["acc"], want, rtol=1e-12, atol=1e-12)
    again = evaluation.evaluate_batch(net, x, L, tgt)["acc"]
    assert torch.equal(res["acc"], again)          # fixed summation order: bit-reproducible


def test_tv_external_upper_incisor_equals_injected(setup):
    evaluation, net, x, L, tgt = setup
    from artspeech_b200.tract_variables import tract_variables_batch
    a = tract_variables_batch(tgt, synthetic.ARTICULATORS, synthetic.UPPER_INCISOR_REF)
    b = tract_variables_batch(synthetic.inject_upper_incisor(tgt), synthetic.TV_ARTICULATORS)
    for u, v in zip(a, b):
        assert torch.equal(u, v)
    with pytest.raises(KeyError):
        tract_variables_batch(tgt, synthetic.ARTICULATORS)   # no upper incisor anywhere


def test_graph_replay_equals_eager(setup):
    evaluation, net, x, L, tgt = setup
    eager = evaluation.evaluate_batch(net, x, L, tgt)
    g = evaluation.GraphedEvaluator(net, x, L, tgt)
    r = g()
    torch.cuda.synchronize()
    assert torch.equal(r["outputs"], eager["outputs"])
    assert torch.equal(r["acc"], eager["acc"])
    # new inputs through the static buffers
    x2 = synthetic.tokens(x.shape[0], x.shape[1], 45, seed=9)
    r2 = g(x2.pin_memory(), None)
    torch.cuda.synchronize()
    want = evaluation.evaluate_batch(net, x2.cuda(), L, tgt)
    assert torch.equal(r2["outputs"], want["outputs"])
    assert torch.equal(r2["acc"], want["acc"])


def test_streamed_evaluator_matches_step_by_step():
    """StreamedEvaluator (copy of step i+1 overlapped with step i) returns, for every step, exactly
    the accumulator vector of the plain step on the same batch."""
    import torch
    from artspeech_b200 import evaluation, models, synthetic
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    net = models.ArtSpeech(45, 10, precision="fast").to(dev).eval()
    B, T = 6, 17
    L = [17, 17, 12, 9, 4, 2]
    batches = []
    for k in range(5):
        tok = synthetic.tokens(B, T, 45, seed=10 + k).pin_memory()
        tgt = synthetic.contours(B * T, 10, 50, seed=20 + k).view(B, T, 10, 2, 50).pin_memory()
        batches.append((tok, tgt))
    want = [evaluation.evaluate_batch(net, t.to(dev), L, g.to(dev))["acc"].cpu() for t, g in batches]
    st = evaluation.StreamedEvaluator(net, batches[0][0].to(dev), L, batches[0][1].to(dev), depth=2)
    got = st.run(batches)
    assert len(got) == 5
    for a, b in zip(got, want):
        assert torch.equal(a, b)
    assert len(st.run(batches[:1])) == 1 and torch.equal(st.run(batches[:1])[0], want[0])


def test_sentence_metrics_match_reference_run_test_numbers():
    """asb_eval_sentence_metrics (per-sentence MED and Pearson correlations, run_test
    evaluation.py:68-84) against numbers produced by the reference's own metrics.py, and the
    run_test info block assembled from the two accumulator vectors."""
    import os
    import numpy as np
    import torch
    from artspeech_b200 import evaluation, metrics, synthetic
    from artspeech_b200.tract_variables import tract_variables_batch
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "sentence_metrics.npz"))
    pred = torch.from_numpy(g["pred"]).cuda()
    tgt = torch.from_numpy(g["target"]).cuda()
    lengths = g["lengths"].tolist()
    ld = torch.tensor(lengths, dtype=torch.int32, device="cuda")
    sacc = evaluation.sentence_metrics_device(pred, tgt, ld)
    want = g["per_sentence"][1:4].sum(axis=1)                      # (3, N): sums over the sentences
    np.testing.assert_allclose(sacc.cpu().numpy(), want, rtol=1e-9, atol=1e-12)
    # the complete per-articulator block of run_test's info dict
    p2cp = metrics.p2cp_distance(pred, tgt)
    tvp = tract_variables_batch(pred, synthetic.ARTICULATORS, synthetic.UPPER_INCISOR_REF.cuda())[0]
    tvt = tract_variables_batch(tgt, synthetic.ARTICULATORS, synthetic.UPPER_INCISOR_REF.cuda())[0]
    acc = evaluation.accumulate_device(p2cp, tvp, tvt, ld)
    info = evaluation.run_test_info(acc, sacc, 10, to_mm=2.0)
    B = len(lengths)
    for i, art in enumerate(synthetic.ARTICULATORS[:10]):
        ref = g["per_sentence"][:, :, i].mean(axis=1)              # means over sentences: p2cp, med, x, y
        assert abs(info[art]["p2cp"] - ref[0]) <= 1e-5 * ref[0]
        assert abs(info[art]["med"] - ref[1]) <= 1e-9 * ref[1]
        assert abs(info[art]["x_corr"] - ref[2]) <= 1e-9
        assert abs(info[art]["y_corr"] - ref[3]) <= 1e-9
        assert info[art]["p2cp_mm"] == info[art]["p2cp"] * 2.0 and info[art]["med_mm"] == info[art]["med"] * 2.0
    # a larger ragged batch against the fp64 oracle
    from oracle import geometry as ogeo
    Bn, T = 37, 23
    t2 = synthetic.contours(Bn * T, 10, 50, seed=31).view(Bn, T, 10, 2, 50)
    p2 = synthetic.perturbed(t2.view(Bn * T, 10, 2, 50), seed=32).view(Bn, T, 10, 2, 50)
    L2 = synthetic.sorted_lengths(Bn, 1, T, seed=33)
    got = evaluation.sentence_metrics_device(p2.cuda(), t2.cuda(), torch.as_tensor(L2, dtype=torch.int32).cuda())
    ref = np.zeros((3, 10))
    for b, L in enumerate([int(v) for v in L2]):
        so, st = p2[b, :L].unsqueeze(0), t2[b, :L].unsqueeze(0)
        ref[0] += ogeo.euclidean_distance(so, st).mean(dim=1)[0].numpy()
        xc, yc = ogeo.pearsons_correlation(so, st)
        ref[1] += xc.mean(dim=-1)[0].numpy()
        ref[2] += yc.mean(dim=-1)[0].numpy()
    np.testing.assert_allclose(got.cpu().numpy(), ref, rtol=1e-8, atol=1e-9)


@pytest.mark.parametrize("precision,tol", [("fp32", 2e-4), ("fast", 2e-2)])
def test_run_test_matches_reference_run_test(golden_dir, tmp_path, precision, tol):
    """The whole ``run_test`` (evaluation.py:17-160) on the drop-in model against the ``info`` dict
    the reference's own run_test returned for the reference model on the same synthetic loader
    (tests/golden/run_test_info.json, written by oracle/make_golden.py), and the files it writes."""
    import json
    import os

    import pandas as pd
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from artspeech_b200 import evaluation, metrics, models
    with open(os.path.join(golden_dir, "run_test_info.json")) as f:
        want = json.load(f)
    torch.manual_seed(0)
    net = models.ArtSpeech(45, 10, precision=precision).cuda()
    info = evaluation.run_test(3, net, synthetic.loader(), metrics.EuclideanDistance("none"), str(tmp_path),
                               list(synthetic.ARTICULATORS))
    assert sorted(info.keys()) == sorted(want.keys())
    assert abs(info["loss"] - want["loss"]) <= tol * want["loss"]
    for art in synthetic.ARTICULATORS:
        for k in ("p2cp", "p2cp_mm", "med", "med_mm"):
            assert abs(info[art][k] - want[art][k]) <= tol * want[art][k], (art, k, info[art][k], want[art][k])
        for k in ("x_corr", "y_corr"):
            assert abs(info[art][k] - want[art][k]) <= 50 * tol, (art, k, info[art][k], want[art][k])
    # files: same names as the reference wrote, TV table with the same columns and close values
    files = sorted(os.listdir(os.path.join(str(tmp_path), "3", "S0_0", "contours")))
    with open(os.path.join(golden_dir, "run_test_S0_0_files.txt")) as f:
        assert files == f.read().split()
    got = pd.read_csv(os.path.join(str(tmp_path), "3", "S1_2", "tract_variables.csv"), dtype={"frame": str})
    ref = pd.read_csv(os.path.join(golden_dir, "run_test_S1_2_tv.csv"), dtype={"frame": str})
    assert list(got.columns) == list(ref.columns) and list(got["frame"]) == list(ref["frame"])
    for name in ("LA", "TTCD", "TBCD", "VEL"):
        a, b = got[f"{name}_target"].to_numpy(), ref[f"{name}_target"].to_numpy()
        assert (abs(a - b) <= 1e-3 * abs(b)).all()            # the reference's fp32 cdist (finding F2)
    ph = pd.read_csv(os.path.join(str(tmp_path), "3", "S1_2", "phonemes.csv"), dtype={"frame": str})
    assert list(ph.columns) == ["sentence", "frame", "phoneme"] and len(ph) == len(got)


def test_two_shards_accumulate_to_the_one_shard_vector(setup):
    """SURVEY.md 8e on hardware: the utterances of a batch split r::2 over two shards, each evaluated on
    its own stream with its own scratch (as two ranks would), and the two `accumulate_device` vectors
    summed (what the all-reduce does) equal the one-shard vector; the per-frame results are bit-equal."""
    evaluation, net, x, L, tgt = setup
    whole = evaluation.evaluate_batch(net, x, L, tgt)
    torch.cuda.synchronize()
    B = x.shape[0]
    total = torch.zeros_like(whole["acc"])
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    parts = []
    for r in range(2):
        idx = evaluation.shard_utterances(B, r, 2)
        Lr = [int(L[i]) for i in idx]                      # still sorted descending
        ws = torch.empty(net.workspace_bytes(len(idx), max(Lr)), dtype=torch.uint8, device="cuda")
        streams[r].wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(streams[r]):
            res = evaluation.evaluate_batch(net, x[idx], Lr, tgt[idx], workspace=ws, overlap_target=False)
        parts.append((idx, Lr, res, ws))
    for s in streams:
        torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    for idx, Lr, res, _ in parts:
        total += res["acc"]
        T_r = res["outputs"].shape[1]
        for k, b in enumerate(idx):
            n = Lr[k]
            assert torch.equal(res["outputs"][k, :n], whole["outputs"][b, :n])
            assert torch.equal(res["p2cp"][k, :n], whole["p2cp"][b, :n])
            assert torch.equal(res["tv_pred"][k, :n], whole["tv_pred"][b, :n])
        assert T_r == max(Lr)
    rel = ((total - whole["acc"]).abs() / whole["acc"].abs().clamp_min(1e-300)).max().item()
    print(f"2-shard accumulators vs 1-shard: max rel {rel:.2e}")
    assert rel <= 1e-12
    n = 10
    assert total[n].item() == B and total[n + 13].item() == float(sum(int(v) for v in L))


def test_graphed_evaluator_follows_weight_changes(setup):
    """A captured step must never run stale weights or freed memory: after an in-place parameter update
    or a load_state_dict the next replay re-captures."""
    evaluation, _, x, L, tgt = setup
    from artspeech_b200 import models
    torch.manual_seed(1)
    net = models.ArtSpeech(45, 10).cuda().eval()
    g = evaluation.GraphedEvaluator(net, x, L, tgt)
    a = g()["outputs"].clone()
    assert not g.stale()
    with torch.no_grad():
        net.linear[0].bias.add_(0.5)                       # bumps the parameter version
    assert g.stale()
    net(x[:2, :5], [5, 5])                                 # an eager call with another shape in between
    b = g()["outputs"].clone()
    torch.cuda.synchronize()
    want = evaluation.evaluate_batch(net, x, L, tgt)["outputs"]
    assert torch.equal(b, want) and (a - b).abs().max().item() > 1e-4
    net.to("cuda")                                         # _apply drops the packed arena
    sd = {k: v.clone() for k, v in net.state_dict().items()}
    net.load_state_dict(sd)
    c = g()["outputs"]
    torch.cuda.synchronize()
    assert torch.equal(c, want)


@pytest.mark.parametrize("precision", ["fast", "fp32"])
def test_repeated_runs_are_bit_identical(precision):
    """Every kernel of the path is deterministic by construction (fixed summation orders, no atomics on results):
    the forward and a whole evaluation with three steps in flight on three streams must repeat bit for bit.
    Stands in for racecheck, which is closed on this pool (tools/determinism_check.py is the long version)."""
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from artspeech_b200 import evaluation, models
    torch.manual_seed(1)
    net = models.ArtSpeech(45, 10, precision=precision).cuda().eval()
    B, T = 41, 37
    x = synthetic.tokens(B, T, 45, seed=4).cuda()
    L = [int(v) for v in synthetic.sorted_lengths(B, 9, T, seed=4)]
    with torch.no_grad():
        first = net(x, L).clone()
        for _ in range(8):
            assert torch.equal(net(x, L), first)
    ld = torch.tensor(L, dtype=torch.int32, device="cuda")
    sets = [(synthetic.tokens(B, T, 45, seed=s).cuda(),
             synthetic.contours(B * T, 10, 50, seed=10 + s, device="cuda").view(B, T, 10, 2, 50)) for s in range(3)]
    pipe = evaluation.PipelinedEvaluator(net, B, T, depth=3)
    totals = []
    for _ in range(4):
        pipe.begin()
        for i in range(7):
            pipe.step(sets[i % 3][0], L, ld, sets[i % 3][1])
        totals.append(pipe.finish(reduce=False).clone())
    torch.cuda.synchronize()
    assert all(torch.equal(t, totals[0]) for t in totals[1:])
    assert float(totals[0][-1]) == 7.0 * sum(L)
