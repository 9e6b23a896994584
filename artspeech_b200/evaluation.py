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


def accumulate_device(p2cp, tv_pred, tv_tgt, lengths_dev, total=None):
    """Same vector as ``accumulate`` computed by the library (asb_eval_accumulate_total): two launches,
    fixed summation order.  All tensors on the GPU; lengths_dev (B) int32.  ``total`` (optional fp64
    vector of the same size on the GPU) also receives the batch's vector ADDED to it: the running sum
    of an evaluation, reduced over ranks once at its end."""
    from . import _lib
    B, T, n = p2cp.shape
    dev = p2cp.device
    scratch = torch.empty(B * (n + 12), dtype=torch.float64, device=dev)
    acc = torch.empty(accumulator_size(n), dtype=torch.float64, device=dev)
    if total is not None and (total.dtype != torch.float64 or total.device != dev or total.numel() != acc.numel()
                              or not total.is_contiguous()):
        raise ValueError("total must be a contiguous float64 tensor of accumulator_size(n) elements on the GPU")
    with torch.cuda.device(dev):
        _lib.check(_lib.lib().asb_eval_accumulate_total(
            _lib.ptr(p2cp.contiguous()), _lib.ptr(tv_pred.contiguous()), _lib.ptr(tv_tgt.contiguous()),
            _lib.ptr(lengths_dev), B, T, n, _lib.ptr(scratch), _lib.ptr(acc),
            _lib.ptr(total) if total is not None else None, _lib.stream_ptr(dev)))
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


class EvalBuffers:
    """Every buffer one evaluation step of a (B, T) batch writes (``asb_eval_buffers``): contours, P2CP,
    tract variables of prediction and target (values, index pairs, points of constriction), the fp64
    accumulator vector, the forward's scratch and - with ``total=True`` - the running sum over the steps
    that use these buffers.  Allocated once and reused by every step of the same shape, so that a step is a
    single native call with no allocation."""

    def __init__(self, model, B, T, total=False, workspace=None):
        from . import _lib
        dev = model.embedding.weight.device
        n, d = model._dims[1], model._dims[4]
        f32 = dict(dtype=torch.float32, device=dev)
        self.shape = (int(B), int(T))
        self.outputs = torch.empty(B, T, n, 2, d, **f32)
        self.p2cp = torch.empty(B, T, n, **f32)
        self.tv_pred, self.tv_target = torch.empty(B, T, 4, **f32), torch.empty(B, T, 4, **f32)
        self.tv_pred_idx = torch.empty(B, T, 4, 2, dtype=torch.int32, device=dev)
        self.tv_target_idx = torch.empty(B, T, 4, 2, dtype=torch.int32, device=dev)
        self.tv_pred_poc, self.tv_target_poc = torch.empty(B, T, 4, 2, 2, **f32), torch.empty(B, T, 4, 2, 2, **f32)
        self.scratch = torch.empty(B * (n + 12), dtype=torch.float64, device=dev)
        self.acc = torch.empty(accumulator_size(n), dtype=torch.float64, device=dev)
        self.total = torch.zeros(accumulator_size(n), dtype=torch.float64, device=dev) if total else None
        need = model.workspace_bytes(B, T)
        if workspace is not None and (workspace.dtype != torch.uint8 or workspace.device != dev
                                      or workspace.numel() < need):
            raise ValueError(f"workspace must be a uint8 tensor of >= {need} bytes on {dev}")
        self.workspace = workspace if workspace is not None else torch.empty(need, dtype=torch.uint8, device=dev)
        c = _lib.AsbEvalBuffers()
        for name in ("outputs", "p2cp", "tv_pred", "tv_pred_idx", "tv_pred_poc", "tv_target", "tv_target_idx",
                     "tv_target_poc", "scratch", "acc", "workspace"):
            setattr(c, name, getattr(self, name).data_ptr())
        c.total = self.total.data_ptr() if self.total is not None else None
        c.workspace_bytes = self.workspace.numel()
        self.c = c

    def result(self):
        return {"outputs": self.outputs, "p2cp": self.p2cp, "tv_pred": self.tv_pred, "tv_target": self.tv_target,
                "tv_pred_idx": self.tv_pred_idx, "tv_target_idx": self.tv_target_idx, "tv_pred_poc": self.tv_pred_poc,
                "tv_target_poc": self.tv_target_poc, "acc": self.acc}


def _tv_rows(articulators, external_upper_incisor):
    import ctypes

    from .tract_variables import _ABI_ORDER, UPPER_INCISOR
    index = articulators if isinstance(articulators, dict) else {name: k for k, name in enumerate(articulators)}
    if external_upper_incisor:
        index = dict(index)
        index.setdefault(UPPER_INCISOR, -1)
    return (ctypes.c_int32 * 6)(*[index[name] for name in _ABI_ORDER])     # KeyError like the reference


@torch.no_grad()
def evaluate_step(model, tokens, lengths, lengths_dev, targets, buffers, articulators=None, upper_incisor=None,
                  side_stream=None):
    """The native step (``asb_eval_step``) on device-resident inputs into preallocated ``EvalBuffers``: ONE
    library call enqueues forward, P2CP, both tract-variable launches and the accumulation on the current
    stream (the target's tract variables on ``side_stream``).  tokens (B, T_in) int64 and targets
    (B, T, N, 2, 50) fp32 contiguous, both on the model's device; ``lengths`` the host list (sorted
    descending, lengths[0] == T == the buffers' T), ``lengths_dev`` the same as an int32 device tensor."""
    import ctypes

    from . import _lib
    from .models import _token_error
    dev = model.embedding.weight.device
    B, T = buffers.shape
    L = [int(v) for v in lengths]
    if len(L) != B or not L or L[0] != T:
        raise ValueError(f"lengths must hold {B} values with lengths[0] == {T}")
    if tokens.device != dev or targets.device != dev or tokens.dtype != torch.int64 or targets.dtype != torch.float32:
        raise ValueError("evaluate_step needs int64 tokens and float32 targets on the model's device")
    if not tokens.is_contiguous() or not targets.is_contiguous() or tuple(targets.shape[:2]) != (B, T):
        raise ValueError(f"targets must be contiguous (B, T, ...) = ({B}, {T}, ...); tokens contiguous")
    articulators = articulators or synthetic.ARTICULATORS
    ui = synthetic.UPPER_INCISOR_REF if upper_incisor is None else upper_incisor
    if ui.device != dev or ui.dtype != torch.float32 or not ui.is_contiguous():
        ui = ui.to(device=dev, dtype=torch.float32).contiguous()
    if ui.shape[-2:] != (2, 50) or ui.numel() not in (100, B * T * 100):
        raise ValueError(f"upper_incisor must be (2, 50) or one (2, 50) contour per frame, got {tuple(ui.shape)}")
    rows = _tv_rows(articulators, True)
    handle = model._ensure_native(dev)
    if not torch.cuda.is_current_stream_capturing():
        _token_error(handle, model._dims[0])
    cur = torch.cuda.current_stream(dev)
    side = cur if side_stream is None else side_stream
    buffers._keep = (ui, tokens, targets, lengths_dev)      # alive until the next step reuses the buffers
    with torch.cuda.device(dev):
        _lib.check(_lib.lib().asb_eval_step(
            handle, _lib.ptr(tokens), tokens.shape[1], (ctypes.c_int32 * B)(*L), _lib.ptr(lengths_dev), B,
            _lib.ptr(targets), rows, _lib.ptr(ui), 0 if ui.numel() == 100 else 100, ctypes.byref(buffers.c),
            ctypes.c_void_p(cur.cuda_stream), ctypes.c_void_p(side.cuda_stream)))
    return buffers.result()


@torch.no_grad()
def evaluate_batch(model, tokens, lengths, targets, articulators=None, upper_incisor=None, mark=None,
                   lengths_dev=None, overlap_target=True, workspace=None, total=None, buffers=None):
    """One evaluation step on the model's device.

    tokens (B, T) int64, lengths sorted descending, targets (B, T, N, 2, 50) fp32, on the GPU (host
    tensors are copied to the model's device first).
    Returns dict(outputs, p2cp (B,T,N), tv_pred (B,T,4), tv_target (B,T,4), acc (N+14,) fp64, and the
    tract variables' index pairs / points of constriction).  The whole step is one native call
    (``evaluate_step``); ``buffers`` (an ``EvalBuffers`` of the right shape) avoids its allocations.
    ``total``: fp64 vector that also receives the accumulators added to it.
    ``mark(name)``, if given, is called before and after the step."""
    mark = mark or (lambda name: None)
    dev = model.embedding.weight.device
    from .models import _host_token_check
    _host_token_check(tokens, model._dims[0])
    tokens = tokens.to(dev, non_blocking=True)        # host batches (pinned) are copied here
    targets = targets.to(dev, non_blocking=True)
    L = [int(v) for v in (lengths.tolist() if torch.is_tensor(lengths) else lengths)]
    if len(L) != tokens.shape[0]:
        raise RuntimeError(f"Expected `len(lengths)` to be equal to batch_size, but got {len(L)} "
                           f"(batch_size={tokens.shape[0]})")
    if lengths_dev is None:
        lengths_dev = torch.as_tensor(L, dtype=torch.int32).to(dev)
    T = max(1, min(L[0], tokens.shape[1])) if L else 1
    tgt = targets[:, :T]
    if not tgt.is_contiguous() or tgt.dtype != torch.float32:
        tgt = tgt.to(torch.float32).contiguous()
    tok = tokens if (tokens.dtype == torch.int64 and tokens.is_contiguous()) else tokens.to(torch.int64).contiguous()
    if buffers is None:
        buffers = EvalBuffers(model, len(L), T, workspace=workspace)
        if total is not None:
            buffers.total = total
            buffers.c.total = total.data_ptr()
    mark("start")
    cur = torch.cuda.current_stream(dev)
    side = _side_stream(dev) if overlap_target else cur
    res = evaluate_step(model, tok, [T] + L[1:] if L and L[0] > T else L, lengths_dev, tgt, buffers, articulators,
                        upper_incisor, side)
    mark("reduce")
    return res


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
    """``evaluate_batch`` for a fixed (batch, T) shape captured once in a CUDA graph and replayed: the
    step is ~10 short kernels, and a replay costs the host one launch.  Inputs are copied into static
    device buffers (``tokens``, ``targets``, ``lengths_dev``); results are the static tensors of the
    captured step (valid until the next replay); ``total`` is the running sum of the accumulator vectors
    of all replays since ``reset_total()``.

    The captured graph holds raw pointers into the model's packed-weight arena and into buffers.  The
    evaluator therefore (1) owns its buffers, scratch included (no eager call can re-allocate them, and
    several evaluators can run side by side), (2) keeps a strong reference to the native handle it
    captured, so the arena cannot be freed under the graph, and (3) remembers the weights key: a replay
    after the weights changed re-captures instead of silently running the old weights."""

    def __init__(self, model, tokens, lengths, targets, articulators=None, upper_incisor=None, warmup=3,
                 overlap_target=True):
        dev = model.embedding.weight.device
        self.model = model
        self.lengths = [int(v) for v in lengths]
        self.T = max(1, min(self.lengths[0], tokens.shape[1])) if self.lengths else 1
        self.lengths_dev = torch.as_tensor(self.lengths, dtype=torch.int32).to(dev)
        self.tokens = tokens.to(dev).to(torch.int64).clone().contiguous()
        self.targets = targets.to(dev)[:, :self.T].to(torch.float32).clone().contiguous()
        self.upper_incisor = (synthetic.UPPER_INCISOR_REF if upper_incisor is None else upper_incisor).to(dev)
        self.upper_incisor = self.upper_incisor.to(torch.float32).contiguous()
        self.articulators = articulators
        self.buffers = EvalBuffers(model, len(self.lengths), self.T, total=True)
        self.total = self.buffers.total
        self.side = torch.cuda.Stream(device=dev) if overlap_target else None

        def run():
            host = [self.T] + self.lengths[1:]
            return evaluate_step(model, self.tokens, host, self.lengths_dev, self.targets, self.buffers,
                                 self.articulators, self.upper_incisor, self.side)

        self._run = run
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):
                run()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self._capture()
        self.reset_total()

    def _capture(self):
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.result = self._run()
        self._handle = self.model._native          # strong reference: the arena the graph points into
        self._key = self.model._weights_key()

    def stale(self):
        """True if the model's weights or packed arena are no longer the captured ones."""
        return self.model._native is not self._handle or self.model._weights_key() != self._key

    def reset_total(self):
        self.total.zero_()

    def set_lengths(self, lengths):
        """New per-utterance lengths for the next replays (sorted descending, none above the captured T):
        the kernels read the lengths from the device vector, so one captured step serves every batch of
        its (B, T) shape; frames at t >= length are computed as padded frames and masked out of the
        accumulators."""
        L = [int(v) for v in lengths]
        if len(L) != len(self.lengths) or L[0] > self.T or min(L) < 1 or any(a < b for a, b in zip(L, L[1:])):
            raise RuntimeError(f"lengths must be {len(self.lengths)} values in [1, {self.T}] sorted in decreasing order")
        self.lengths = L
        self.lengths_dev.copy_(torch.as_tensor(L, dtype=torch.int32), non_blocking=True)

    def replay(self):
        if self.stale():
            # weights changed (optimizer step, load_state_dict, .to()): capture again with the current ones
            # (one eager step first, on a side stream, so that the weights are packed outside the capture)
            dev = self.tokens.device
            keep = self.total.clone()
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):
                self._run()
            torch.cuda.current_stream(dev).wait_stream(side)
            torch.cuda.synchronize(dev)
            self._capture()
            self.total.copy_(keep)
        self.graph.replay()
        return self.result

    def __call__(self, tokens=None, targets=None, lengths=None):
        if tokens is not None:
            self.tokens.copy_(tokens, non_blocking=True)
        if targets is not None:
            self.targets.copy_(targets if targets.shape[1] == self.T else targets[:, :self.T], non_blocking=True)
        if lengths is not None:
            self.set_lengths(lengths)
        return self.replay()


class PipelinedEvaluator:
    """Evaluation of device-resident batches with ``depth`` steps in flight on ``depth`` streams.

    A step is a chain of kernels that do not fill the machine one by one (the two recurrent layers run
    2 x B/16 CTAs for T dependent time steps; the geometry kernels are fp32-pipe bound): with the next
    step's chain on another stream the recurrent kernels of step i+1 run beside the geometry kernels of
    step i.  Every slot owns its buffers and its running total; ``finish()`` joins the streams, adds
    the slots' totals and sums them over ranks with ONE all-reduce per evaluation (run_test aggregates
    once, evaluation.py:142-159).  Measured at batch 256 x 100 (tools/pipeline_depth.py, 60 steps): 0.437 / 0.372 /
    0.351 / 0.351 ms per step at depth 1 / 2 / 3 / 4."""

    def __init__(self, model, batch, seq_len, articulators=None, upper_incisor=None, depth=3):
        dev = model.embedding.weight.device
        self.model, self.device, self.depth = model, dev, depth
        self.articulators = articulators
        ui = synthetic.UPPER_INCISOR_REF if upper_incisor is None else upper_incisor
        self.upper_incisor = ui.to(device=dev, dtype=torch.float32).contiguous()
        self.buffers = [EvalBuffers(model, batch, seq_len, total=True) for _ in range(depth)]
        least, greatest = torch.cuda.Stream.priority_range()
        self.streams = [torch.cuda.Stream(device=dev, priority=greatest) for _ in range(depth)]
        self.sides = [torch.cuda.Stream(device=dev, priority=least) for _ in range(depth)]
        self.step_index = 0
        self._fork = torch.cuda.Event()

    def begin(self):
        """Zero the running totals and order the slots' streams after the current stream."""
        cur = torch.cuda.current_stream(self.device)
        for b in self.buffers:
            b.total.zero_()
        self._fork.record(cur)
        for s in self.streams:
            s.wait_event(self._fork)
        self.step_index = 0

    def step(self, tokens, lengths, lengths_dev, targets):
        """Enqueue one batch on the next slot's stream (returns at once)."""
        k = self.step_index % self.depth
        self.step_index += 1
        with torch.cuda.stream(self.streams[k]):
            return evaluate_step(self.model, tokens, lengths, lengths_dev, targets, self.buffers[k],
                                 self.articulators, self.upper_incisor, self.sides[k])

    def finish(self, reduce=True):
        """Join the slots into the current stream; returns the evaluation's accumulator vector (summed over
        the slots and, with ``reduce``, over the ranks - one all-reduce)."""
        cur = torch.cuda.current_stream(self.device)
        for s in self.streams:
            cur.wait_stream(s)
        total = self.buffers[0].total.clone()
        for b in self.buffers[1:]:
            total += b.total
        return reduce_accumulators(total) if reduce else total


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
        # masked mean over the valid frames (:59-63), kept on the device: one host sync per evaluation, not per batch
        if loss.dim() >= 2:
            m = mask.view(*mask.shape, *([1] * (loss.dim() - 2))).to(loss.dtype)
            losses.append((loss * m).sum() / (m.sum() * max(1, loss[0, 0].numel())))
        else:
            losses.append(loss.detach().reshape(()))
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
    info = {"loss": float(torch.stack([v.double() for v in losses]).mean().item()) if losses else float("nan")}
    to_mm = dataset_config.RES * dataset_config.PIXEL_SPACING                  # :150
    info.update(run_test_info(acc, sacc, n, to_mm=to_mm, articulators=articulators))
    return info
