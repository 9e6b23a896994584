"""Device-resident evaluation step: forward -> P2CP -> tract variables -> per-utterance means.

This is the batched form of the inner loops of the reference's ``run_test``
(phoneme_to_articulation/encoder_decoder/evaluation.py:53-128) and of ``tract_variables``
(phoneme_to_articulation/__init__.py:201-297): the reference moves the outputs to the CPU and
loops per sentence / per frame in Python; here everything up to the small accumulator vector
stays on the GPU.  Utterances are independent, so a multi-GPU run shards them by rank and sums
the accumulators with one all-reduce (``reduce_accumulators``).
"""
from __future__ import annotations

import torch

from . import synthetic
from .metrics import p2cp_distance
from .tract_variables import tract_variables_batch

# layout of the fp64 accumulator vector (all sums over utterances / valid frames of this rank)
#   [0:N)        sum over utterances of the per-utterance mean P2CP, per articulator (evaluation.py:72,82)
#   [N]          number of utterances
#   [N+1 : N+5)  sum over valid frames of the predicted TVs (LA, TTCD, TBCD, VEL)
#   [N+5 : N+9)  same for the target TVs
#   [N+9 : N+13) sum over valid frames of |TV_pred - TV_target|
#   [N+13]       number of valid frames


def accumulator_size(n_articulators):
    return n_articulators + 14


def length_mask(lengths, T, device):
    L = torch.as_tensor(lengths, device=device, dtype=torch.int64).view(-1, 1)
    return torch.arange(T, device=device).view(1, -1) < L  # (B, T), helpers.py:79-91 make_padding_mask


def accumulate(p2cp, tv_pred, tv_tgt, lengths):
    """Per-frame results -> the fp64 accumulator vector (layout above).  Pure tensor ops on the
    inputs' device; p2cp (B,T,N), tv_* (B,T,4), lengths (B)."""
    B, T, n = p2cp.shape
    m = length_mask(lengths, T, p2cp.device).unsqueeze(-1).to(torch.float64)
    Lf = m.sum(dim=1)                                                 # (B, 1)
    per_utt = (p2cp.to(torch.float64) * m).sum(dim=1) / Lf            # (B, N) mean over the valid frames
    acc = torch.zeros(accumulator_size(n), dtype=torch.float64, device=p2cp.device)
    acc[:n] = per_utt.sum(dim=0)
    acc[n] = float(B)
    acc[n + 1:n + 5] = (tv_pred.to(torch.float64) * m).sum(dim=(0, 1))
    acc[n + 5:n + 9] = (tv_tgt.to(torch.float64) * m).sum(dim=(0, 1))
    acc[n + 9:n + 13] = ((tv_pred - tv_tgt).abs().to(torch.float64) * m).sum(dim=(0, 1))
    acc[n + 13] = m.sum()
    return acc


def accumulate_device(p2cp, tv_pred, tv_tgt, lengths_dev):
    """Same vector as ``accumulate`` computed by the library (asb_eval_accumulate): two launches,
    fixed summation order.  All tensors on the GPU; lengths_dev (B) int32."""
    from . import _lib
    B, T, n = p2cp.shape
    dev = p2cp.device
    scratch = torch.empty(B * (n + 12), dtype=torch.float64, device=dev)
    acc = torch.empty(accumulator_size(n), dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        _lib.check(_lib.lib().asb_eval_accumulate(
            _lib.ptr(p2cp.contiguous()), _lib.ptr(tv_pred.contiguous()), _lib.ptr(tv_tgt.contiguous()),
            _lib.ptr(lengths_dev), B, T, n, _lib.ptr(scratch), _lib.ptr(acc), _lib.stream_ptr(dev)))
    return acc


def sentence_metrics_device(outputs, targets, lengths_dev):
    """Sums over the batch's sentences of the per-sentence MED and Pearson correlations per
    articulator (asb_eval_sentence_metrics; run_test, evaluation.py:68-84): (3, N) fp64 on the GPU,
    rows = MED, x_corr, y_corr.  outputs/targets (B,T,N,2,D) fp32 on the GPU, lengths_dev (B) int32."""
    from . import _lib
    B, T, n, _, D = outputs.shape
    dev = outputs.device
    tgt = targets[:, :T]
    scratch = torch.empty(B * n * 3, dtype=torch.float64, device=dev)
    acc = torch.empty(3 * n, dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        _lib.check(_lib.lib().asb_eval_sentence_metrics(
            _lib.ptr(outputs.contiguous()), _lib.ptr(tgt.contiguous()), _lib.ptr(lengths_dev), B, T, n, D,
            _lib.ptr(scratch), _lib.ptr(acc), _lib.stream_ptr(dev)))
    return acc.view(3, n)


def run_test_info(acc, sentence_acc, n_articulators, to_mm=1.0, articulators=None):
    """The per-articulator block of run_test's ``info`` dict (evaluation.py:146-159) from the reduced
    accumulators: {art: {x_corr, y_corr, p2cp, p2cp_mm, med, med_mm}}."""
    articulators = articulators or synthetic.ARTICULATORS
    a = acc.detach().cpu().tolist()
    s = sentence_acc.detach().cpu().view(3, -1).tolist()
    n = n_articulators
    utt = a[n]
    info = {}
    for i, art in enumerate(articulators[:n]):
        p2cp, med = a[i] / utt, s[0][i] / utt
        info[art] = {"x_corr": s[1][i] / utt, "y_corr": s[2][i] / utt, "p2cp": p2cp, "p2cp_mm": p2cp * to_mm,
                     "med": med, "med_mm": med * to_mm}
    return info


def shard_utterances(n_utterances, rank, world):
    """Indices of the utterances rank ``rank`` of ``world`` owns (round-robin)."""
    return list(range(rank, n_utterances, world))


_SIDE_STREAMS = {}


def _side_stream(dev):
    key = (dev.type, dev.index if dev.index is not None else torch.cuda.current_device())
    if key not in _SIDE_STREAMS:
        _SIDE_STREAMS[key] = torch.cuda.Stream(device=dev)
    return _SIDE_STREAMS[key]


@torch.no_grad()
def evaluate_batch(model, tokens, lengths, targets, articulators=None, upper_incisor=None, mark=None,
                   lengths_dev=None, overlap_target=True, workspace=None):
    """One evaluation step on the model's device.

    tokens (B, T) int64, lengths sorted descending, targets (B, T, N, 2, 50) fp32, on the GPU (host
    tensors are copied to the model's device first).
    Returns dict(outputs, p2cp (B,T,N), tv_pred (B,T,4), tv_target (B,T,4), acc (N+14,) fp64).
    ``mark(name)``, if given, is called at the stage boundaries (bench.py records CUDA events there).
    """
    mark = mark or (lambda name: None)
    articulators = articulators or synthetic.ARTICULATORS
    if upper_incisor is None:
        upper_incisor = synthetic.UPPER_INCISOR_REF
    dev = model.embedding.weight.device
    tokens = tokens.to(dev, non_blocking=True)        # host batches (pinned) are copied here
    targets = targets.to(dev, non_blocking=True)
    if lengths_dev is None:
        lengths_dev = torch.as_tensor([int(v) for v in lengths], dtype=torch.int32).to(dev)
    mark("start")
    # The target's tract variables do not depend on the forward pass: they run on a side stream
    # (forked here, joined before the accumulation) and fill the SMs the recurrent kernels leave
    # idle (2 directions x B/16 CTAs).  Inside a CUDA-graph capture this becomes a parallel branch.
    cur = torch.cuda.current_stream(dev)
    side = _side_stream(dev) if overlap_target else cur
    T_in = min(int(max(lengths)), tokens.shape[1]) if len(lengths) else 0
    tgt = targets[:, :max(1, T_in)]
    if not tgt.is_contiguous():
        tgt = tgt.contiguous()
    fork = torch.cuda.Event()
    fork.record(cur)
    outputs = model(tokens, lengths, lengths_dev=lengths_dev, workspace=workspace)    # evaluation.py:57
    # The branch is enqueued AFTER the forward's kernels (it depends only on `fork`): a full wave of
    # tract-variable CTAs takes every register of every SM, so the recurrent kernel (B/16 CTAs per
    # direction, the critical path) has to be placed first.  Measured on B200 at batch 256 x 100
    # (tools/overlap_check.py): 0.500 -> 0.479 ms per step; capping the branch's grid
    # (tract_variables_batch(max_ctas=...)) to leave SMs free did not improve on that.
    side.wait_event(fork)
    with torch.cuda.stream(side):
        tv_tgt, _, _ = tract_variables_batch(tgt, articulators, upper_incisor)    # __init__.py:261
    mark("forward")
    assert outputs.shape[1] == tgt.shape[1], (outputs.shape, tgt.shape)
    p2cp = p2cp_distance(outputs, tgt)                                # evaluation.py:72 (all sentences at once)
    mark("p2cp")
    tv_pred, _, _ = tract_variables_batch(outputs, articulators, upper_incisor)   # __init__.py:256
    cur.wait_stream(side)
    if side is not cur and not torch.cuda.is_current_stream_capturing():
        tv_tgt.record_stream(cur)
    mark("tv")
    acc = accumulate_device(p2cp, tv_pred, tv_tgt, lengths_dev)       # evaluation.py:68-84,142-156
    mark("reduce")
    return {"outputs": outputs, "p2cp": p2cp, "tv_pred": tv_pred, "tv_target": tv_tgt, "acc": acc}


def reduce_accumulators(acc):
    """Sum the accumulator vector over ranks (NCCL all-reduce when torch.distributed is up)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(acc, op=dist.ReduceOp.SUM)
    return acc


def summarize(acc, n_articulators, articulators=None):
    """Accumulators -> the numbers run_test reports (evaluation.py:142-159), in normalised units."""
    articulators = articulators or synthetic.ARTICULATORS
    a = acc.detach().cpu().tolist()
    n = n_articulators
    utt, frames = a[n], a[n + 13]
    out = {art: {"p2cp": a[i] / utt} for i, art in enumerate(articulators[:n])}
    for k, name in enumerate(("LA", "TTCD", "TBCD", "VEL")):
        out[name] = {"pred_mean": a[n + 1 + k] / frames, "target_mean": a[n + 5 + k] / frames,
                     "abs_err_mean": a[n + 9 + k] / frames}
    out["utterances"] = utt
    out["frames"] = frames
    return out


class GraphedEvaluator:
    """``evaluate_batch`` for a fixed (batch, lengths) shape captured once in a CUDA graph and replayed:
    the step is ~10 short kernels, so eager launches leave the GPU idle between them.  Inputs are
    copied into static device buffers (``tokens``, ``targets``); results are the static tensors of
    the captured step (valid until the next replay)."""

    def __init__(self, model, tokens, lengths, targets, articulators=None, upper_incisor=None, warmup=3,
                 overlap_target=True):
        dev = model.embedding.weight.device
        self.model = model
        # The captured graph holds raw pointers into the model's packed-weight arena and into a scratch
        # buffer.  The evaluator therefore (1) owns its scratch buffer (no eager call can re-allocate it,
        # and several evaluators can run side by side), (2) keeps a strong reference to the native handle
        # it captured, so the arena cannot be freed under the graph, and (3) remembers the weights key: a
        # replay after the weights changed re-captures instead of silently running the old weights.
        self.lengths = [int(v) for v in lengths]
        T_run = max(1, min(self.lengths[0], tokens.shape[1])) if self.lengths else 1
        self.workspace = torch.empty(model.workspace_bytes(len(self.lengths), T_run), dtype=torch.uint8, device=dev)
        self.lengths_dev = torch.as_tensor(self.lengths, dtype=torch.int32).to(dev)
        self.tokens = tokens.to(dev).clone()
        self.targets = targets.to(dev).clone()
        self.upper_incisor = (synthetic.UPPER_INCISOR_REF if upper_incisor is None else upper_incisor).to(dev)
        self.articulators = articulators

        def run():
            return evaluate_batch(model, self.tokens, self.lengths, self.targets, self.articulators,
                                  self.upper_incisor, lengths_dev=self.lengths_dev, overlap_target=overlap_target,
                                  workspace=self.workspace)

        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):
                run()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self._run = run
        self._capture()

    def _capture(self):
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.result = self._run()
        self._handle = self.model._native          # strong reference: the arena the graph points into
        self._key = self.model._weights_key()

    def stale(self):
        """True if the model's weights or packed arena are no longer the captured ones."""
        return self.model._native is not self._handle or self.model._weights_key() != self._key

    def replay(self):
        if self.stale():
            # weights changed (optimizer step, load_state_dict, .to()): capture again with the current ones
            # (one eager step first, on a side stream, so that the weights are packed outside the capture)
            dev = self.tokens.device
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):
                self._run()
            torch.cuda.current_stream(dev).wait_stream(side)
            torch.cuda.synchronize(dev)
            self._capture()
        self.graph.replay()
        return self.result

    def __call__(self, tokens=None, targets=None):
        if tokens is not None:
            self.tokens.copy_(tokens, non_blocking=True)
        if targets is not None:
            self.targets.copy_(targets, non_blocking=True)
        return self.replay()


class StreamedEvaluator:
    """Evaluation of a SEQUENCE of host batches with the copies overlapped with the compute.

    The reference's ``run_test`` loop (evaluation.py:53-128) takes one batch at a time from the
    DataLoader, moves it to the device, runs it and moves the results back.  Here ``depth`` captured
    steps (``GraphedEvaluator``) alternate: while step i runs, the host->device copy of step i+1
    (tokens + target contours from pinned memory) proceeds on a copy stream, and the accumulator
    vector of each step comes back with an asynchronous device->host copy into pinned memory.
    Every step still copies its own inputs in and its own result out; only the waiting is hidden.
    """

    def __init__(self, model, tokens, lengths, targets, articulators=None, upper_incisor=None, depth=2):
        dev = model.embedding.weight.device
        self.device = dev
        self.depth = depth
        self.slots = [GraphedEvaluator(model, tokens, lengths, targets, articulators, upper_incisor)
                      for _ in range(depth)]
        n = self.slots[0].result["acc"].numel()
        self.acc_host = [torch.empty(n, dtype=torch.float64).pin_memory() for _ in range(depth)]
        self.copy_stream = torch.cuda.Stream(device=dev)
        self.copied = [torch.cuda.Event() for _ in range(depth)]
        self.done = [torch.cuda.Event() for _ in range(depth)]

    def run(self, batches):
        """batches: iterable of (tokens, targets) pinned host tensors of the captured shape.
        Returns the list of per-step accumulator vectors (fp64 CPU tensors, summed over ranks)."""
        out = []
        cur = torch.cuda.current_stream(self.device)
        n_steps = 0
        for i, (tok, tgt) in enumerate(batches):
            s = i % self.depth
            slot = self.slots[s]
            if i >= self.depth:
                self.done[s].synchronize()        # step i - depth finished: its slot and host buffer are free
                out.append(self.acc_host[s].clone())
            with torch.cuda.stream(self.copy_stream):
                slot.tokens.copy_(tok, non_blocking=True)
                slot.targets.copy_(tgt, non_blocking=True)
                self.copied[s].record(self.copy_stream)
            cur.wait_event(self.copied[s])
            slot.replay()
            reduce_accumulators(slot.result["acc"])
            self.acc_host[s].copy_(slot.result["acc"], non_blocking=True)
            self.done[s].record(cur)
            n_steps = i + 1
        for i in range(max(0, n_steps - self.depth), n_steps):
            s = i % self.depth
            self.done[s].synchronize()
            out.append(self.acc_host[s].clone())
        return out


def save_outputs(sentences_ids, frame_ids, outputs, targets, lengths, phonemes, articulators, save_to,
                 regularize_out=False):
    """phoneme_to_articulation/__init__.py:143-199 with the same arguments: per sentence,
    ``contours/{frame}_{art}[_true].npy`` for the valid frames and ``phonemes.csv``.  B-spline
    regularisation (``vt_tools.bs_regularization``, un-vendored) is not restated."""
    import os

    import pandas as pd

    from .contour_files import save_contours
    if regularize_out:
        raise NotImplementedError("regularize_out needs vt_tools.bs_regularization.regularize_Bsplines, "
                                  "which is not part of the reference repository")
    out_h, tgt_h = outputs.detach().cpu(), targets.detach().cpu()
    for b, (sentence_id, length, sentence_phonemes, sentence_frames) in enumerate(
            zip(sentences_ids, lengths, phonemes, frame_ids)):
        n = min(int(length), len(sentence_phonemes), len(sentence_frames))
        saves_sentence_dir = os.path.join(save_to, sentence_id)
        save_contours(saves_sentence_dir, list(sentence_frames)[:n], out_h[b, :n], tgt_h[b, :n], articulators)
        rows = [{"sentence": sentence_id, "frame": f, "phoneme": p}
                for f, p in zip(list(sentence_frames)[:n], list(sentence_phonemes)[:n])]
        pd.DataFrame(rows).to_csv(os.path.join(saves_sentence_dir, "phonemes.csv"), index=False)


@torch.no_grad()
def run_test(epoch, model, dataloader, criterion, outputs_dir, articulators, device=None, regularize_out=False,
             save=True):
    """``run_test`` of the reference (phoneme_to_articulation/encoder_decoder/evaluation.py:17-160)
    with the same arguments and the same ``info`` dict, for the drop-in model.

    The reference moves every batch's outputs to the CPU and loops per sentence (P2CP, MED, Pearson)
    and per frame (tract variables) in Python.  Here the per-sentence metrics are the device kernels
    ``asb_p2cp`` + ``asb_eval_accumulate`` and ``asb_eval_sentence_metrics`` over the whole batch,
    the tract variables of all frames are two launches (``tract_variables`` driver), and only the
    accumulators and the files to be written leave the GPU.  ``save=False`` (extension) skips the
    per-frame ``.npy`` / CSV files.  Batches are the dataset's 8-tuples (dataset.py collate):
    (sentences_ids, sentences, targets, lengths, phonemes, reference_arrays, sentence_frames, voicing)."""
    import os

    from . import _lib
    from .tract_variables import REQUIRED_ARTICULATORS_FOR_TVS, UPPER_INCISOR
    from .tract_variables import tract_variables as tv_driver
    if device is None:
        device = torch.device("cuda")
    device = torch.device(device)
    if device.type != "cuda":
        raise RuntimeError("artspeech_b200.evaluation.run_test runs on CUDA only; there is no CPU fallback")
    _lib.lib()
    epoch_outputs_dir = os.path.join(outputs_dir, str(epoch))
    os.makedirs(epoch_outputs_dir, exist_ok=True)
    model.eval()
    dataset_config = dataloader.dataset.dataset_config
    articulators = list(articulators)
    n = len(articulators)
    acc = torch.zeros(accumulator_size(n), dtype=torch.float64, device=device)
    sacc = torch.zeros(3, n, dtype=torch.float64, device=device)
    losses = []
    for (sentences_ids, sentences, targets, lengths, phonemes, reference_arrays, sentence_frames, _) in dataloader:
        sentences = sentences.to(device)
        targets = targets.to(device)
        L = [int(v) for v in (lengths.tolist() if torch.is_tensor(lengths) else lengths)]
        lengths_dev = torch.as_tensor(L, dtype=torch.int32).to(device)
        outputs = model(sentences, L)                                          # :57
        T = outputs.shape[1]
        tgt = targets[:, :T].contiguous()
        loss = criterion(outputs, tgt)                                         # :58-63
        mask = length_mask(L, T, device)
        losses.append(loss[mask].mean().item() if loss.dim() >= 2 else float(loss))
        p2cp = p2cp_distance(outputs, tgt)                                     # :65-84, all sentences at once
        zeros4 = torch.zeros(outputs.shape[0], T, 4, dtype=torch.float32, device=device)
        acc += accumulate_device(p2cp, zeros4, zeros4, lengths_dev)
        sacc += sentence_metrics_device(outputs, tgt, lengths_dev)
        if UPPER_INCISOR not in articulators:                                  # :93-109
            tv_articulators = sorted(articulators + [UPPER_INCISOR])
            ref_idx = tv_articulators.index(UPPER_INCISOR)
            ref = reference_arrays.to(device=device, dtype=outputs.dtype)[:, :T]
            outputs = torch.concat([outputs[:, :, :ref_idx], ref, outputs[:, :, ref_idx:]], dim=2)
            tgt = torch.concat([tgt[:, :, :ref_idx], ref, tgt[:, :, ref_idx:]], dim=2)
        else:
            tv_articulators = articulators
        if save:
            if all(a in tv_articulators for a in REQUIRED_ARTICULATORS_FOR_TVS):   # :113-129
                tv_driver(sentences_ids, sentence_frames, outputs, tgt, L, phonemes, tv_articulators,
                          epoch_outputs_dir)
            save_outputs(sentences_ids, sentence_frames, outputs, tgt, L, phonemes, tv_articulators,
                         epoch_outputs_dir, regularize_out)                    # :131-141
    reduce_accumulators(acc)
    reduce_accumulators(sacc)
    info = {"loss": float(sum(losses) / len(losses)) if losses else float("nan")}
    to_mm = dataset_config.RES * dataset_config.PIXEL_SPACING                  # :150
    info.update(run_test_info(acc, sacc, n, to_mm=to_mm, articulators=articulators))
    return info
