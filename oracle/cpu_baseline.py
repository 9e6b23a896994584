"""Timed CPU baseline: the reference's evaluation path restated with the reference's own CPU ops.

TEST / BENCH INFRASTRUCTURE ONLY (see ``oracle/__init__.py``); used by ``bench.py``'s
``cpu_baseline`` leg and ``--impl reference`` arm.  The reference checkout does not travel to the
GPU box, so this "port" reproduces what the reference executes on the host, op for op:

* forward: stock torch CPU ``nn.Embedding`` / ``nn.GRU`` (packed) / ``nn.Linear`` / ``nn.LayerNorm``
  (``oracle.model.artspeech_forward``; models.py:126-145), fp32;
* P2CP per sentence with ``torch.cdist`` + two ``min`` (phoneme_to_articulation/metrics.py:41-44,
  called as evaluation.py:68-72);
* tract variables per frame, for prediction and target, in a Python loop with ``torch.cdist``,
  two ``min`` and three ``.item()`` per search (tract_variables.py:23-35 driven by
  phoneme_to_articulation/__init__.py:242-261), after injecting the upper incisor
  (evaluation.py:93-109).
"""
from __future__ import annotations

import time

import torch

from artspeech_b200 import synthetic
from oracle import geometry as og
from oracle import model as omodel


def _p2cp_cdist(u, v):
    """metrics.py:32-46 with reduction 'none' (fp32, torch.cdist as the reference)."""
    n, m = u.shape[-2], v.shape[-2]
    dist = torch.cdist(u, v)
    u2cp, _ = dist.min(axis=-1)
    v2cp, _ = dist.min(axis=-2)
    return (torch.sum(u2cp, dim=-1) / n + torch.sum(v2cp, dim=-1) / m) / 2


def _tv_cdist(arr1, arr2):
    """tract_variables.py:23-35."""
    d = torch.cdist(arr1, arr2)
    min_d0, argmin_d0 = d.min(dim=0)
    min_d1, argmin_d1 = min_d0.min(dim=0)
    i = argmin_d0[argmin_d1].item()
    j = argmin_d1.item()
    return min_d1.item(), arr1[i], arr2[j]


def _frame_tvs(frame):
    """tract_variables.py:73-125 for one (11, 2, 50) frame given as name -> (50, 2) views."""
    d = {n: frame[k].T for k, n in enumerate(synthetic.TV_ARTICULATORS)}
    s = og.ART_SLICES
    tongue, uinc, spm = d[og.TONGUE], d[og.UPPER_INCISOR], d[og.SOFT_PALATE_MIDLINE]
    la = _tv_cdist(d[og.LOWER_LIP], d[og.UPPER_LIP])
    ttcd = _tv_cdist(tongue[slice(*s["tongue-tip"])], uinc[slice(*s["upper-incisor"])])
    palate = torch.cat([uinc[slice(*s["hard-palate"])], spm[slice(*s["soft-palate"])]], axis=0)
    tbcd = _tv_cdist(tongue[slice(*s["tongue-body"])], palate)
    vel = _tv_cdist(spm[slice(*s["velum"])], d[og.PHARYNX])
    return la[0], ttcd[0], tbcd[0], vel[0]


@torch.no_grad()
def evaluate_batch_cpu(sd, tokens, lengths, targets):
    """One evaluation step the way the reference runs it on the host.  Returns
    (per-utterance mean P2CP (B, N), TV sums pred (4), TV sums target (4), frames)."""
    outputs = omodel.artspeech_forward(sd, tokens, lengths, torch.float32)         # evaluation.py:57
    T = outputs.shape[1]
    targets = targets[:, :T]
    per_utt = []
    for so, st, L in zip(outputs, targets, lengths):                               # evaluation.py:68-72
        L = int(L)
        p = _p2cp_cdist(so[:L].unsqueeze(0).transpose(-1, -2), st[:L].unsqueeze(0).transpose(-1, -2))
        per_utt.append(p.mean(dim=1)[0])
    out11 = synthetic.inject_upper_incisor(outputs)                                # evaluation.py:93-109
    tgt11 = synthetic.inject_upper_incisor(targets)
    tv_p = [0.0] * 4
    tv_t = [0.0] * 4
    frames = 0
    for so, st, L in zip(out11, tgt11, lengths):                                   # __init__.py:222-261
        for f in range(int(L)):
            a = _frame_tvs(so[f])
            b = _frame_tvs(st[f])
            for k in range(4):
                tv_p[k] += a[k]
                tv_t[k] += b[k]
            frames += 1
    return torch.stack(per_utt), tv_p, tv_t, frames


class ReferencePath:
    """The reference's OWN code for the path, loaded by file path from the checkout or from the copies
    staged under oracle/_ref/ (oracle/stage_ref.py): ``ArtSpeech`` (encoder_decoder/models.py:99-145),
    ``p2cp_distance`` / ``euclidean_distance`` / ``pearsons_correlation`` (metrics.py:9-68) and
    ``calculate_vocal_tract_variables`` (tract_variables.py:73-125), driven exactly like run_test's inner
    loops (encoder_decoder/evaluation.py:53-84) and the frame driver
    (phoneme_to_articulation/__init__.py:222-261)."""

    def __init__(self, sd, vocab=45, n_art=10):
        from oracle import reference_loader as ref
        if not ref.baseline_available():
            raise RuntimeError("neither /root/reference nor oracle/_ref is present")
        self.metrics = ref.root_metrics()
        self.tv = ref.tract_variables().calculate_vocal_tract_variables
        self.net = ref.models().ArtSpeech(vocab, n_art).eval()
        self.net.load_state_dict(sd)

    @torch.no_grad()
    def evaluate_batch(self, tokens, lengths, targets):
        m = self.metrics
        outputs = self.net(tokens, lengths)                                         # evaluation.py:57
        targets = targets[:, :outputs.shape[1]]
        per_utt = []
        for so, st, L in zip(outputs, targets, lengths):                            # evaluation.py:68-84
            so, st = so[:int(L)].unsqueeze(dim=0), st[:int(L)].unsqueeze(dim=0)
            p2cp = m.p2cp_distance(so, st).mean(dim=1)
            m.euclidean_distance(so, st).mean(dim=1)
            x_corr, y_corr = m.pearsons_correlation(so, st)
            x_corr.mean(dim=-1)[0], y_corr.mean(dim=-1)[0]
            per_utt.append(p2cp[0])
        out11 = synthetic.inject_upper_incisor(outputs)                             # evaluation.py:93-109
        tgt11 = synthetic.inject_upper_incisor(targets)
        names = synthetic.TV_ARTICULATORS
        tv_p, tv_t, frames = [0.0] * 4, [0.0] * 4, 0
        for so, st, L in zip(out11, tgt11, lengths):                                # __init__.py:222-261
            for out, target in zip(so[:int(L)], st[:int(L)]):
                a = self.tv({art: tensor.T for art, tensor in zip(names, out)})
                b = self.tv({art: tensor.T for art, tensor in zip(names, target)})
                for k, name in enumerate(("LA", "TTCD", "TBCD", "VEL")):
                    tv_p[k] += a[name]["value"]
                    tv_t[k] += b[name]["value"]
                frames += 1
        return torch.stack(per_utt), tv_p, tv_t, frames


def make_evaluator(sd, prefer_reference=True):
    """(kind, evaluate(tokens, lengths, targets)): the reference's own files when they are present
    ("reference"), else the op-for-op port above ("port")."""
    if prefer_reference:
        try:
            path = ReferencePath(sd)
            return "reference", path.evaluate_batch
        except Exception:
            pass
    return "port", lambda tokens, lengths, targets: evaluate_batch_cpu(sd, tokens, lengths, targets)


def best_thread_count(evaluate, cores):
    """The path is thousands of tiny torch ops (per-frame cdist / min / .item()): more intra-op threads can
    make it SLOWER.  Try a few counts on a small sample and keep the fastest, so that the CPU gets its best."""
    tokens = synthetic.tokens(1, 24, 45, seed=3)
    targets = synthetic.contours(24, 10, 50, seed=4).view(1, 24, 10, 2, 50)
    best, best_t = None, cores
    for t in sorted({cores, max(1, cores // 2), min(cores, 4), 1}, reverse=True):
        torch.set_num_threads(t)
        evaluate(tokens[:, :4], [4], targets[:, :4])
        t0 = time.perf_counter()
        evaluate(tokens, [24], targets)
        dt = time.perf_counter() - t0
        if best is None or dt < best:
            best, best_t = dt, t
    torch.set_num_threads(best_t)
    return best_t


def time_cpu_baseline(sd, batch, seq_len, repeats=1, seed=0, threads=None, prefer_reference=True):
    """Wall-clock frames/s of the CPU evaluation of (batch, seq_len) synthetic utterances.
    Returns (frames/s, seconds, kind)."""
    kind, evaluate = make_evaluator(sd, prefer_reference)
    if threads:
        torch.set_num_threads(threads)
    tokens = synthetic.tokens(batch, seq_len, 45, seed=seed)
    lengths = [seq_len] * batch
    targets = synthetic.contours(batch * seq_len, 10, 50, seed=seed + 1).view(batch, seq_len, 10, 2, 50)
    evaluate(tokens[:1, :8], [8], targets[:1, :8])  # warm-up of the op dispatch paths
    best = None
    for _ in range(repeats):
        t0 = time.perf_counter()
        evaluate(tokens, lengths, targets)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return batch * seq_len / best, best, kind
