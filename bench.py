"""Headline benchmark: articulatory frames/sec of the evaluation hot path (forward + P2CP + TVs).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --steps K --warmup W    # the reference's CPU path (its own files)

Workload = BASELINE.json configs[1]: the encoder_decoder model at batch 256, sequence length 100,
followed by P2CP and the tract variables of prediction and target for every frame.  One "step" is
one such batch (25 600 frames).  Under torchrun every rank runs its own batches (utterance sharding,
weak scaling); the fp64 metric accumulators are summed on the device over the steps and reduced over
the ranks with ONE NCCL all-reduce at the end of the timed evaluation (run_test aggregates once,
encoder_decoder/evaluation.py:142-159).

Prints ONE JSON line (rank 0).
  value     frames/s with inputs resident in HBM: K steps through PipelinedEvaluator (each step ONE
            native call, asb_eval_step; three steps in flight on three streams, two input sets of 103 MB
            alternate, per-step working set 335 MB > L2), the WHOLE region between two CUDA events,
            the all-reduce included, max over ranks.  `serial` holds the round-1 measure (one captured
            step at a time, L2 flushed between steps, per-step events) beside it.
  e2e       the same steps through the public API from pinned HOST buffers (H2D of tokens + targets and
            D2H of the accumulators inside the timed region), wall clock.
  roofline  dominant kernel of the step; kernel_ms are per-kernel device times measured with the launches
            queued AHEAD of the device (so no host gap is inside an interval).
  fp32_mode the same device-resident measure in fp32 precision (the 1e-5 mode).
  gpu_baseline  the unchanged reference module graph under stock torch on the same GPU (cuDNN GRU, cuBLAS
            linears, torch.cdist) - the existing-kernel bar on this box.
  configs   BASELINE.json configs[2..4]: 1 M-frame geometry, 64 x 1024 recurrence, 65 536-utterance sweep.
  cpu_baseline  the reference's own CPU code (oracle/_ref, staged copies of its files) on a bounded sample.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "articulatory frames/sec (fwd + P2CP/TV)"
UNIT = "frames/s"
VOCAB, N_ART, D = 45, 10, 50
# per-frame work of the model
MAC_DECODER = 32768 + 10 * 123904       # linear.0 + ten predictors
MAC_PROJ_L1 = 2 * 384 * 256
MAC_RECURRENT = 196608                  # the sequential h-path, both layers and directions
BYTES_PROJ_L1 = 4096                    # out0 row (256 fp32) read + gi row (768 fp32) written
BYTES_P2CP = 8040                       # 2*10*400 read + 40 written
BYTES_TV = 5024                         # pred + target: 2 * (6*400 + 4*28)
BYTES_AREA = 2404                       # 2 walls * 50 * 8 read + (2*200*4 + 4) written


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return {"hbm_gbs": p["hbm_gbs"], "tflops_burst": p["bf16_tflops"],
                "tflops_sustained": p.get("bf16_tflops_sustained", p["bf16_tflops"]), "source": "measured"}
    # fallback stated in /opt/skills/guides/B200_PROFILING.md
    return {"hbm_gbs": 6650.0, "tflops_burst": 1590.0, "tflops_sustained": 1400.0, "source": "fallback"}


def workload_config(world, B, T, precision):
    """The `config` both arms print (the reference arm times a bounded sample of the same workload)."""
    return {"workload": f"encoder_decoder BiGRU fwd + P2CP + TVs (pred+target), batch {B} x len {T} per GPU "
                        f"(configs[1]), 10 articulators x 50 points",
            "precision": precision, "frames_per_step_per_gpu": B * T,
            "parallelism": f"utterance-sharded x{world}; accumulators summed on the device over the steps, ONE NCCL "
                           f"all-reduce of {N_ART + 14} fp64 per evaluation",
            "l2": "two input sets (103 MB each) alternate, per-step working set 335 MB > L2; the `serial` record "
                  "flushes 256 MiB between steps",
            "launch": "each step is one native call (asb_eval_step); three steps in flight on three streams "
                      "(PipelinedEvaluator); the whole K-step region is timed between two CUDA events"}


class ClockSampler:
    """SM clock + throttle reasons sampled every ms through NVML while the timed region runs (NVML is set up
    before the region; the region itself adds one sample while its queue is still draining)."""

    def __init__(self, index):
        self.index = index
        self.samples = []
        self.reasons = set()
        self.max_mhz = None
        self._stop = threading.Event()
        self.thread = None
        self.error = None

    def _init(self):
        """NVML set-up, done synchronously in start(): it takes longer than a short timed region."""
        import pynvml as nv
        nv.nvmlInit()
        self._nv = nv
        self._h = nv.nvmlDeviceGetHandleByIndex(self.index)
        self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(self._h, nv.NVML_CLOCK_SM))
        self._flags = {
            "hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown",
                                   getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0)),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown",
                                           getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0)),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown",
                                           getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0)),
            "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap",
                                    getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0)),
        }
        self._get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons",
                                    getattr(nv, "nvmlDeviceGetCurrentClocksThrottleReasons", None))

    def sample(self):
        self.samples.append(float(self._nv.nvmlDeviceGetClockInfo(self._h, self._nv.NVML_CLOCK_SM)))
        if self._get_reasons is not None:
            r = int(self._get_reasons(self._h))
            for name, bit in self._flags.items():
                if bit and (r & int(bit)):
                    self.reasons.add(name)

    def _run(self):
        try:
            while not self._stop.is_set():
                self.sample()
                self._stop.wait(0.001)
        except Exception as exc:  # report it instead of inventing numbers
            self.error = repr(exc)

    def start(self):
        try:
            self._init()
        except Exception as exc:  # NVML missing
            self.error = repr(exc)
            return
        self.thread = threading.Thread(target=self._run, daemon=True)
        self.thread.start()

    def stop(self):
        self._stop.set()
        if self.thread is not None:
            self.thread.join(timeout=2)
        sm = sorted(self.samples)
        out = {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.max_mhz, "samples": len(sm),
               "reasons": sorted(self.reasons)}
        if self.error:
            out["error"] = self.error
        return out


def visible_index(local_rank):
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        ids = [v for v in vis.split(",") if v != ""]
        if local_rank < len(ids) and ids[local_rank].isdigit():
            return int(ids[local_rank])
    return local_rank


def bind_near_gpu(index):
    """Pin this process to the CPUs NVML reports as local to GPU `index`, so that the pinned host
    buffers of the end-to-end leg are allocated on the GPU's own NUMA node (with several ranks on one
    host, buffers on a remote node halve the host->device copy rate).  Returns (old mask, #cpus)."""
    try:
        import pynvml as nv
        nv.nvmlInit()
        h = nv.nvmlDeviceGetHandleByIndex(index)
        words = nv.nvmlDeviceGetCpuAffinity(h, ((os.cpu_count() or 64) + 63) // 64)
        near = {64 * i + b for i, w in enumerate(words) for b in range(64) if (int(w) >> b) & 1}
        old = os.sched_getaffinity(0)
        use = near & old
        if use and use != old:
            os.sched_setaffinity(0, use)
            return old, len(use)
        return None, len(old)
    except Exception:
        return None, None


# ------------------------------------------------------------------------------------------ reference arm
def run_reference(args):
    """The reference's own CPU implementation of the path on the box's host cores: its files staged under
    oracle/_ref (oracle/stage_ref.py; /root/reference does not exist on the GPU box), all host threads, each
    step a bounded sample of the configs[1] batch."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import cpu_baseline as cb
    from oracle import model as omodel
    from artspeech_b200 import synthetic
    world = int(os.environ.get("WORLD_SIZE", "1"))
    sd = omodel.make_state_dict(VOCAB, N_ART, seed=0)
    kind, evaluate = cb.make_evaluator(sd)
    # all host threads it can USE: the count (of those available) at which this path runs fastest
    cores = cb.best_thread_count(evaluate, len(os.sched_getaffinity(0)) or os.cpu_count() or 1)
    sb, sl = args.ref_batch, args.seq_len
    tokens = synthetic.tokens(sb, sl, VOCAB, seed=0)
    lengths = [sl] * sb
    targets = synthetic.contours(sb * sl, N_ART, D, seed=1).view(sb, sl, N_ART, 2, D)
    for _ in range(max(args.warmup, 1) if args.warmup > 0 else 0):
        evaluate(tokens[:2], lengths[:2], targets[:2])
    t0 = time.perf_counter()
    for _ in range(args.steps):
        evaluate(tokens, lengths, targets)
    dt = time.perf_counter() - t0
    value = args.steps * sb * sl / dt
    sample = (f"{sb} of the batch's {args.batch} utterances x {sl} frames per step (utterances are independent: the "
              f"per-frame rate does not depend on the batch size), fp32, torch CPU ops, "
              + ("the reference's own models.py / metrics.py / tract_variables.py (oracle/_ref)" if kind == "reference"
                 else "oracle port (oracle/_ref not staged)"))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(world, args.batch, sl, args.precision),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------ stock torch on the GPU
def stock_p2cp(outputs, targets):
    """phoneme_to_articulation/metrics.py:32-46 via metrics.py:38-52, as written, on CUDA tensors."""
    u, v = outputs.transpose(-1, -2), targets.transpose(-1, -2)
    n, m = u.shape[-2], v.shape[-2]
    dist = torch.cdist(u, v)
    u2cp, _ = dist.min(axis=-1)
    v2cp, _ = dist.min(axis=-2)
    return (torch.sum(u2cp, dim=-1) / n + torch.sum(v2cp, dim=-1) / m) / 2


def stock_tvs(frames11, index):
    """tract_variables.py:23-125 batched over frames with stock torch ops (cdist + two min + gathers)."""
    def tv(a1, a2):                                   # (F, 2, P), (F, 2, R)
        d = torch.cdist(a1.transpose(1, 2), a2.transpose(1, 2))     # (F, P, R)
        min_d0, arg0 = d.min(dim=1)
        val, j = min_d0.min(dim=1)
        i = arg0.gather(1, j[:, None])[:, 0]
        return val, i, j
    g = lambda name: frames11[:, index[name]]
    tongue, uinc, spm = g("tongue"), g("upper-incisor"), g("soft-palate-midline")
    out = [tv(g("lower-lip"), g("upper-lip")), tv(tongue[:, :, 30:45], uinc[:, :, 25:50]),
           tv(tongue[:, :, 10:30], torch.cat([uinc[:, :, 0:25], spm[:, :, 35:50]], dim=2)),
           tv(spm[:, :, 0:15], g("pharynx"))]
    return torch.stack([o[0] for o in out], dim=1)


def measure_gpu_baseline(model, tokens_d, lengths, targets_d, flush, steps):
    """The unchanged reference module graph (models.py:126-145: nn.Embedding, cuDNN nn.GRU on a packed
    sequence, cuBLAS linears, LayerNorm, stack, sigmoid = ArtSpeech._library_graph over the same parameters)
    and the reference's metric formulas under stock torch on this GPU; CUDA events, L2 flushed between."""
    from artspeech_b200 import synthetic

    def timed(fn, n):
        for _ in range(2):
            fn()
        pairs = []
        for _ in range(n):
            flush.zero_()
            s_ = torch.cuda.Event(enable_timing=True); e_ = torch.cuda.Event(enable_timing=True)
            s_.record(); fn(); e_.record()
            pairs.append((s_, e_))
        torch.cuda.synchronize()
        return sum(a.elapsed_time(b) for a, b in pairs) / len(pairs)

    n = max(3, min(steps, 10))
    out = {"kind": "stock-torch", "torch": torch.__version__, "cudnn": torch.backends.cudnn.version()}
    with torch.no_grad():
        keep_mm, keep_cd = torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32
        try:
            fwd = lambda: model._library_graph(tokens_d, lengths)
            out["fwd_ms"] = timed(fwd, n)                                   # torch defaults (fp32 matmul, cuDNN tf32)
            torch.backends.cuda.matmul.allow_tf32 = True
            torch.backends.cudnn.allow_tf32 = True
            out["fwd_tf32_ms"] = timed(fwd, n)
            outputs = fwd()
        finally:
            torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32 = keep_mm, keep_cd
        out["p2cp_ms"] = timed(lambda: stock_p2cp(outputs, targets_d), n)
        o11 = synthetic.inject_upper_incisor(outputs).flatten(0, 1)
        t11 = synthetic.inject_upper_incisor(targets_d).flatten(0, 1)
        out["tv_ms"] = timed(lambda: (stock_tvs(o11, synthetic.TV_ART_INDEX), stock_tvs(t11, synthetic.TV_ART_INDEX)), n)
    frames = tokens_d.shape[0] * tokens_d.shape[1]
    total = out["fwd_tf32_ms"] + out["p2cp_ms"] + out["tv_ms"]
    out["step_ms"] = total
    out["frames_per_s"] = frames / (total * 1e-3)
    out["note"] = ("forward = ArtSpeech._library_graph (the reference's module graph over the same parameters); step = "
                   "fwd_tf32 + p2cp + tv, kernels back to back; the reference's own TV loop (3 .item() per search, "
                   "0.35 ms per frame) is not timed - the batched cdist form is the favourable stock version")
    return out


# ------------------------------------------------------------------------------------------ our arm
def run_ours(args):
    import torch.distributed as dist
    from artspeech_b200 import _lib, evaluation, models, synthetic
    from artspeech_b200 import area_function as af
    from artspeech_b200.metrics import p2cp_distance
    from artspeech_b200.tract_variables import tract_variables_batch

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: artspeech_b200 has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    _lib.lib()  # fail loudly here if the CUDA library is missing
    peaks = load_peaks()
    old_affinity, near_cpus = bind_near_gpu(visible_index(local_rank)) if world > 1 else (None, None)

    B, T, K = args.batch, args.seq_len, args.steps
    W = max(args.warmup, 3)
    frames = B * T
    n_acc = N_ART + 14
    torch.manual_seed(0)                       # same initial weights as the reference constructor
    model = models.ArtSpeech(VOCAB, N_ART, precision=args.precision).to(dev).eval()
    # synthetic utterances of this rank (sharded by utterance: rank r owns its own batches); two input sets
    tokens_h = synthetic.tokens(B, T, VOCAB, seed=100 + rank).pin_memory()
    targets_h = synthetic.contours(frames, N_ART, D, seed=200 + rank).view(B, T, N_ART, 2, D).pin_memory()
    lengths = [T] * B
    lengths_dev = torch.tensor(lengths, dtype=torch.int32, device=dev)
    tok_sets = [tokens_h.to(dev), synthetic.tokens(B, T, VOCAB, seed=300 + rank).to(dev)]
    tgt_sets = [targets_h.to(dev),
                synthetic.contours(frames, N_ART, D, seed=400 + rank, device=dev).view(B, T, N_ART, 2, D)]
    tokens_d, targets_d = tok_sets[0], tgt_sets[0]
    uinc_d = synthetic.UPPER_INCISOR_REF.to(dev)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()

    def max_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def pipelined_region(net, steps, warm, sampler=None):
        """K steps, three in flight; ONE all-reduce of the running totals at the end; returns (ms, total)."""
        pipe = evaluation.PipelinedEvaluator(net, B, T, upper_incisor=uinc_d, depth=3)
        pipe.begin()
        for i in range(warm):
            pipe.step(tok_sets[i & 1], lengths, lengths_dev, tgt_sets[i & 1])
        pipe.finish(reduce=True)
        torch.cuda.synchronize(); barrier()
        s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
        l0 = _lib.launch_count()
        s.record()
        pipe.begin()
        for i in range(steps):
            pipe.step(tok_sets[i & 1], lengths, lengths_dev, tgt_sets[i & 1])
        total = pipe.finish(reduce=True)
        e.record()
        if sampler is not None and sampler.thread is not None:
            sampler.sample()                    # the queue is still draining: one sample is certainly under load
        torch.cuda.synchronize()
        launches = _lib.launch_count() - l0
        barrier()
        return max_over_ranks(s.elapsed_time(e)), total, launches

    # ---------------- device-resident timing (value)
    pipelined_region(model, 2, W)              # first-touch, NCCL communicator, clocks
    sampler = ClockSampler(visible_index(local_rank))
    sampler.start()
    total_ms, total_acc, launches = pipelined_region(model, K, W, sampler)
    clocks = sampler.stop()
    value = world * frames * K / (total_ms * 1e-3)
    summary = evaluation.summarize(total_acc, N_ART)
    # every rank's frames and utterances arrived in the single all-reduce
    assert summary["frames"] == float(world * frames * K), (summary["frames"], world * frames * K)
    assert summary["utterances"] == float(world * B * K), summary["utterances"]

    # ---------------- the round-1 measure: one captured step at a time, L2 flushed between steps
    graphed = evaluation.GraphedEvaluator(model, tokens_d, lengths, targets_d, upper_incisor=uinc_d)
    for _ in range(W):
        graphed()
    torch.cuda.synchronize()
    pairs = []
    for i in range(K):
        flush.zero_()                                   # evict L2 (not timed)
        s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
        s.record(); graphed(); e.record()
        pairs.append((s, e))
    torch.cuda.synchronize()
    serial_ms = max_over_ranks(sum(a.elapsed_time(b) for a, b in pairs) / K)
    serial = {"ms_per_step": serial_ms, "value": world * frames / (serial_ms * 1e-3),
              "how": "CUDA-graph replay of one captured step, 256 MiB L2 flush between steps (not timed), per-step "
                     "CUDA events (BENCH_r01's measure)"}

    # ---------------- per-kernel device times, launches queued ahead of the device
    # A spin kernel (torch.cuda._sleep) runs while the host enqueues the stage's launches and event
    # records, so every interval between two events holds kernel time only - no host launch gap.
    spin = int(250e-6 * 1.9e9)
    model.profile(True)
    model.read_profile(reset=True)
    with torch.no_grad():
        for _ in range(K):
            flush.zero_()
            torch.cuda._sleep(spin)
            model(tokens_d, lengths, lengths_dev=lengths_dev)
    torch.cuda.synchronize()
    prof = model.read_profile(reset=True)
    model.profile(False)
    kernel_ms = {k: (v[0] / v[1] if v[1] else 0.0) for k, v in prof.items()}
    with torch.no_grad():
        outputs_d = model(tokens_d, lengths)
    p2cp_d = p2cp_distance(outputs_d, targets_d)
    tvp_d = tract_variables_batch(outputs_d, synthetic.ARTICULATORS, uinc_d)[0]
    tvt_d = tract_variables_batch(targets_d, synthetic.ARTICULATORS, uinc_d)[0]

    def time_calls(fn, n=None):
        for _ in range(3):
            fn()
        pairs_ = []
        for _ in range(n or K):
            flush.zero_()
            torch.cuda._sleep(spin // 4)
            s_ = torch.cuda.Event(enable_timing=True); e_ = torch.cuda.Event(enable_timing=True)
            s_.record(); fn(); e_.record()
            pairs_.append((s_, e_))
        torch.cuda.synchronize()
        return sum(a_.elapsed_time(b_) for a_, b_ in pairs_) / len(pairs_)

    kernel_ms.update({
        "p2cp": time_calls(lambda: p2cp_distance(outputs_d, targets_d)),
        "tv_pred": time_calls(lambda: tract_variables_batch(outputs_d, synthetic.ARTICULATORS, uinc_d)),
        "tv_target": time_calls(lambda: tract_variables_batch(targets_d, synthetic.ARTICULATORS, uinc_d)),
        "accumulate": time_calls(lambda: evaluation.accumulate_device(p2cp_d, tvp_d, tvt_d, lengths_dev))})
    kernel_sum = sum(kernel_ms.values())

    # ---------------- roofline of the dominant kernel
    kernels = {
        "decoder": ("tensor", 2.0 * MAC_DECODER * frames, kernel_ms.get("decoder", 0.0)),
        "gru_l0": ("tensor", 2.0 * MAC_RECURRENT / 2 * frames, kernel_ms.get("gru_l0", 0.0)),
        "gru_l1": ("tensor", 2.0 * MAC_RECURRENT / 2 * frames, kernel_ms.get("gru_l1", 0.0)),
        # 96 FLOP per byte, below the machine balance (~210): the projection is bound by its 1 KB read + 3 KB written per frame
        "proj_l1": ("hbm", float(BYTES_PROJ_L1) * frames, kernel_ms.get("proj_l1", 0.0)),
        "p2cp": ("hbm", float(BYTES_P2CP) * frames, kernel_ms.get("p2cp", 0.0)),
        "tract_variables": ("hbm", float(BYTES_TV) * frames, kernel_ms["tv_pred"] + kernel_ms["tv_target"]),
    }
    top = max(kernels, key=lambda k: kernels[k][2])
    bound, work, ms = kernels[top]
    if bound == "tensor":
        achieved = work / (ms * 1e-3) / 1e12 if ms > 0 else 0.0
        peak, unit = peaks["tflops_sustained"], "TFLOP/s"
    else:
        achieved = work / (ms * 1e-3) / 1e9 if ms > 0 else 0.0
        peak, unit = peaks["hbm_gbs"], "GB/s"
    traffic, traffic_source = None, None
    tpath = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if os.path.exists(tpath) and (B, T) == (256, 100):
        with open(tpath) as f:
            tj = json.load(f)
        traffic = tj.get("bytes_per_launch", {}).get(top)
        traffic_source = (f"profiles/ncu_traffic.json ({tj.get('capture', 'ncu --set full capture')}): "
                          "dram__bytes_read.sum + dram__bytes_write.sum per launch; NOT measured in this run")
    frac_of = lambda v: ((v[1] / (v[2] * 1e-3) / (1e12 if v[0] == "tensor" else 1e9)) /
                         (peaks["tflops_sustained"] if v[0] == "tensor" else peaks["hbm_gbs"])) if v[2] > 0 else None
    roofline = {"kernel": top, "bound": bound, "achieved": achieved, "peak": peak, "unit": unit,
                "frac": achieved / peak if peak else None,
                "traffic": traffic, "traffic_source": traffic_source,
                "peak_source": f"{peaks['source']} (dense bf16 cuBLAS sustained; the decoder's GEMMs are tcgen05 kind::f16 "
                               f"with fp16 operands, the same tensor rate)" if bound == "tensor" else peaks["source"],
                "ms_per_launch": ms,
                "all_kernels_ms": {k: round(v[2], 4) for k, v in kernels.items()},
                "all_kernels_frac": {k: (round(frac_of(v), 4) if frac_of(v) is not None else None)
                                     for k, v in kernels.items()}}

    # ---------------- fp32 precision (the 1e-5 mode), same device-resident measure
    fp32_mode = None
    if not args.no_fp32 and args.precision != "fp32":
        model32 = models.ArtSpeech(VOCAB, N_ART, precision="fp32")
        model32.load_state_dict(model.state_dict())
        model32 = model32.to(dev).eval()
        k32 = max(2, K)
        ms32, acc32, _ = pipelined_region(model32, k32, 3)
        s32 = evaluation.summarize(acc32, N_ART)
        fp32_mode = {"value": world * frames * k32 / (ms32 * 1e-3), "unit": UNIT, "ms_per_step": ms32 / k32,
                     "steps": k32, "dtype": "f32 accumulate over split fp16 operands (hi + lo, 22 significand bits, tcgen05); contours within 1e-5 of the fp32 reference",
                     "mean_p2cp_first_articulator": s32[synthetic.ARTICULATORS[0]]["p2cp"]}
        del model32

    # ---------------- stock torch on the same GPU (rank 0 of a single-GPU run)
    gpu_baseline = None
    if world == 1 and not args.no_gpu_baseline:
        gpu_baseline = measure_gpu_baseline(model, tokens_d, lengths, targets_d, flush, K)
        gpu_baseline["ours_over_stock"] = {
            "forward": gpu_baseline["fwd_tf32_ms"] / sum(kernel_ms[k] for k in ("gather", "gru_l0", "proj_l1", "gru_l1", "decoder")),
            "p2cp": gpu_baseline["p2cp_ms"] / kernel_ms["p2cp"],
            "tv": gpu_baseline["tv_ms"] / (kernel_ms["tv_pred"] + kernel_ms["tv_target"]),
            "step": gpu_baseline["step_ms"] / (total_ms / K)}

    # ---------------- BASELINE.json configs[2..4]
    configs = {}
    if not args.no_configs:
        if world == 1:
            configs["c3_geometry_1M"] = config3_geometry(dev, peaks, args.c3_frames)
            configs["c4_64x1024"] = config4_recurrence(model, dev)
        configs["c5_65536"] = config5_sweep(model, dev, rank, world, args.c5_utterances, max_over_ranks, barrier)

    # ---------------- end to end from pinned host buffers (e2e)
    # Every step copies its tokens + targets from pinned host memory and reads its accumulator vector
    # back; StreamedEvaluator overlaps the copy of step i+1 with the compute of step i.
    streamed = evaluation.StreamedEvaluator(model, tokens_d, lengths, targets_d, upper_incisor=uinc_d, depth=3)
    r = graphed(tokens_h, targets_h)
    torch.cuda.synchronize()
    want_acc = r["acc"].cpu()
    streamed.run([(tokens_h, targets_h)] * 3)
    barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    accs_h = streamed.run([(tokens_h, targets_h)] * K)
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    if world == 1:
        assert len(accs_h) == K and torch.equal(accs_h[-1], want_acc), "streamed result differs"
    e2e_value = world * frames * K / e2e_s
    h2d = tokens_h.numel() * tokens_h.element_size() + targets_h.numel() * targets_h.element_size()
    d2h = want_acc.numel() * want_acc.element_size()
    # the bare host->device copy of the same bytes, all ranks at once: what the host fabric gives
    copy_stream = torch.cuda.Stream(device=dev)
    barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    with torch.cuda.stream(copy_stream):
        for _ in range(K):
            tok_sets[1].copy_(tokens_h, non_blocking=True)
            tgt_sets[1].copy_(targets_h, non_blocking=True)
    copy_stream.synchronize()
    bare_s = max_over_ranks(time.perf_counter() - t0)
    h2d_gbs_bare = h2d * K / bare_s / 1e9
    h2d_gbs_pipeline = h2d * K / e2e_s / 1e9

    if old_affinity is not None:
        os.sched_setaffinity(0, old_affinity)

    # ---------------- CPU baseline (rank 0, single-GPU run only): the reference's own code on the host cores
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import cpu_baseline as cb
        sd = {k: v.detach().cpu() for k, v in model.state_dict().items()}
        cores = cb.best_thread_count(cb.make_evaluator(sd)[1], len(os.sched_getaffinity(0)) or os.cpu_count() or 1)
        v, secs, kind = cb.time_cpu_baseline(sd, args.cpu_batch, T, repeats=1, threads=cores)
        cpu_baseline = {"value": v, "unit": UNIT, "cores": cores, "kind": kind,
                        "sample": f"{args.cpu_batch} utterances x {T} frames of the same workload ({secs:.1f} s of host "
                                  f"time), fp32 torch CPU ops, per-sentence metrics and per-frame TV loop as run_test; "
                                  + ("the reference's own files staged under oracle/_ref" if kind == "reference"
                                     else "oracle port (oracle/_ref not staged)")}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": total_ms / K, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None,
            "dtype": "fp16 tcgen05 operands, fp32 accumulate and state; fp32 geometry" if args.precision == "fast" else "f32 accumulate over split fp16 operands (hi + lo); fp32 geometry",
            "data": "synthetic",
            "config": workload_config(world, B, T, args.precision),
            "serial": serial,
            "kernel_ms": {k: round(v, 4) for k, v in kernel_ms.items()},
            "kernel_ms_sum": round(kernel_sum, 4),
            "kernel_ms_how": "CUDA events with the launches queued behind a spin kernel (no host gap inside an interval), "
                             "L2 flushed before each; sum <= serial ms_per_step x 1.1 is the consistency check, the "
                             "pipelined ms_per_step is below the sum because three steps overlap",
            "roofline": roofline, "fp32_mode": fp32_mode, "gpu_baseline": gpu_baseline, "configs": configs,
            "cpu_baseline": cpu_baseline,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "pipeline": "depth-3: the H2D of steps i+1, i+2 on a copy stream overlaps step i",
                    "h2d_gbs_pipeline": h2d_gbs_pipeline, "h2d_gbs_bare": h2d_gbs_bare,
                    "frac_of_bare_copy": h2d_gbs_pipeline / h2d_gbs_bare if h2d_gbs_bare else None,
                    "host_cpus_near_gpu": near_cpus},
            "gpu_launches": int(launches), "clocks": clocks,
            "check": {"mean_p2cp_first_articulator": summary[synthetic.ARTICULATORS[0]]["p2cp"],
                      "frames_accumulated": summary["frames"], "frames_expected": float(world * frames * K),
                      "utterances_accumulated": summary["utterances"], "allreduces_in_timed_region": 1 if world > 1 else 0},
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def _time_kernel(fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(iters + 1)]
    ev[0].record()
    for i in range(iters):
        fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    ts = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(iters))
    return ts[len(ts) // 2]


def config3_geometry(dev, peaks, F):
    """configs[2]: P2CP + tract variables + area function over F distinct synthetic frames (inputs >> L2)."""
    from artspeech_b200 import area_function as af
    from artspeech_b200 import synthetic
    from artspeech_b200.metrics import p2cp_distance
    from artspeech_b200.tract_variables import tract_variables_batch
    chunk = 131072
    tgt = torch.empty(F, N_ART, 2, D, dtype=torch.float32, device=dev)
    for k, lo in enumerate(range(0, F, chunk)):
        n = min(chunk, F - lo)
        tgt[lo:lo + n] = synthetic.contours(n, N_ART, D, seed=5000 + k, device=dev)
    pred = synthetic.perturbed(tgt, seed=7)
    uinc = synthetic.UPPER_INCISOR_REF.to(dev)
    hbm = peaks["hbm_gbs"]
    out = {"frames": F, "inputs": "F distinct frames generated on the device (arcs + noise; prediction = target + "
                                  "N(0, 0.01^2)); median of 5 launches, inputs far larger than L2"}
    ms = _time_kernel(lambda: p2cp_distance(pred, tgt))
    out["p2cp"] = {"ms": ms, "frames_per_s": F / ms * 1e3, "gbs": F * BYTES_P2CP / ms * 1e-6,
                   "frac": F * BYTES_P2CP / ms * 1e-6 / hbm}
    ms = _time_kernel(lambda: (tract_variables_batch(pred, synthetic.ARTICULATORS, uinc),
                               tract_variables_batch(tgt, synthetic.ARTICULATORS, uinc)))
    out["tv"] = {"ms": ms, "frames_per_s": F / ms * 1e3, "gbs": F * BYTES_TV / ms * 1e-6,
                 "frac": F * BYTES_TV / ms * 1e-6 / hbm}
    del pred, tgt
    wi = torch.empty(F, 50, 2, dtype=torch.float32, device=dev)
    we = torch.empty(F, 50, 2, dtype=torch.float32, device=dev)
    for k, lo in enumerate(range(0, F, chunk)):
        n = min(chunk, F - lo)
        a, b = synthetic.walls(n, seed=6000 + k, device=dev)
        wi[lo:lo + n], we[lo:lo + n] = a, b
    grid = af.DeviceGrid(af.build_semipolar_grid(**synthetic.AREA_GRID_ARGS), dev)
    ms = _time_kernel(lambda: af.area_function_batch(wi, we, grid))
    out["area"] = {"ms": ms, "frames_per_s": F / ms * 1e3, "gbs": F * BYTES_AREA / ms * 1e-6,
                   "frac": F * BYTES_AREA / ms * 1e-6 / hbm}
    return out


def config4_recurrence(model, dev):
    """configs[3]: batch 64 x length 1024, 1024 dependent recurrent steps per direction and layer."""
    from artspeech_b200 import synthetic
    Bc, Tc = 64, 1024
    x = synthetic.tokens(Bc, Tc, VOCAB, seed=1).to(dev)
    L = [Tc] * Bc
    ld = torch.tensor(L, dtype=torch.int32, device=dev)
    spin = int(250e-6 * 1.9e9)
    with torch.no_grad():
        for _ in range(2):
            model(x, L, lengths_dev=ld)
        model.profile(True)
        model.read_profile(reset=True)
        for _ in range(5):
            torch.cuda._sleep(spin)
            model(x, L, lengths_dev=ld)
        torch.cuda.synchronize()
        prof = model.read_profile(reset=True)
        model.profile(False)
    ms = {k: v[0] / v[1] for k, v in prof.items() if v[1]}
    fwd = sum(ms.values())
    return {"batch": Bc, "seq_len": Tc, "kernel_ms": {k: round(v, 4) for k, v in ms.items()},
            "us_per_step_layer0": ms["gru_l0"] * 1e3 / Tc, "us_per_step_layer1": ms["gru_l1"] * 1e3 / Tc,
            "forward_ms": fwd, "frames_per_s_forward": Bc * Tc / (fwd * 1e-3)}


def config5_sweep(model, dev, rank, world, n_utt, max_over_ranks, barrier):
    """configs[4]: n_utt synthetic utterances, lengths U[50, 100] sorted inside each batch of 256, rank r
    owns batches r::world (utterance sharding), every batch resident in HBM; forward + P2CP + TVs +
    accumulation per batch through PipelinedEvaluator (one native call per batch, variable lengths on the
    device) and ONE all-reduce of the accumulators at the end.  STRONG scaling: the total work is fixed."""
    import torch.distributed as dist
    from artspeech_b200 import evaluation, synthetic
    Bc, Tc = 256, 100
    n_batches = n_utt // Bc
    mine = list(range(rank, n_batches, world))
    tok = torch.empty(len(mine), Bc, Tc, dtype=torch.int64, device=dev)
    tgt = torch.empty(len(mine), Bc, Tc, N_ART, 2, D, dtype=torch.float32, device=dev)
    host_len, dev_len, valid = [], [], 0
    for k, b in enumerate(mine):
        tok[k] = synthetic.tokens(Bc, Tc, VOCAB, seed=9000 + b).to(dev)
        tgt[k] = synthetic.contours(Bc * Tc, N_ART, D, seed=20000 + b, device=dev).view(Bc, Tc, N_ART, 2, D)
        L = synthetic.sorted_lengths(Bc, 50, Tc, seed=b)
        host_len.append([int(v) for v in L])
        dev_len.append(L.to(dev))
        valid += int(L.sum())
    uinc = synthetic.UPPER_INCISOR_REF.to(dev)
    pipe = evaluation.PipelinedEvaluator(model, Bc, Tc, upper_incisor=uinc, depth=3)

    def run():
        pipe.begin()
        for k in range(len(mine)):
            pipe.step(tok[k], host_len[k], dev_len[k], tgt[k])
        return pipe.finish(reduce=True)

    run()
    torch.cuda.synchronize(); barrier()
    s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
    s.record()
    total = run()
    e.record()
    torch.cuda.synchronize()
    ms = max_over_ranks(s.elapsed_time(e))
    summ = evaluation.summarize(total, N_ART)
    all_valid = torch.tensor([float(valid)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(all_valid)
    assert summ["frames"] == float(all_valid.item()) and summ["utterances"] == float(n_batches * Bc), summ
    return {"utterances": n_batches * Bc, "valid_frames": summ["frames"], "padded_frames": n_batches * Bc * Tc,
            "n_gpus": world, "ms": ms, "frames_per_s": summ["frames"] / (ms * 1e-3),
            "utterances_per_s": n_batches * Bc / (ms * 1e-3), "scaling": "strong", "allreduces": 1 if world > 1 else 0,
            "mean_p2cp_first_articulator": summ[synthetic.ARTICULATORS[0]]["p2cp"],
            "lengths": "U[50,100] sorted inside each batch of 256; padded frames are computed and masked out"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--seq-len", type=int, default=100)
    ap.add_argument("--precision", default="fast", choices=["fast", "fp32"])
    ap.add_argument("--cpu-batch", type=int, default=256,
                    help="utterances in the bounded CPU-baseline sample (256 = one whole configs[1] batch, ~10 s of host time)")
    ap.add_argument("--ref-batch", type=int, default=16, help="utterances per step of --impl reference")
    ap.add_argument("--c3-frames", type=int, default=1048576)
    ap.add_argument("--c5-utterances", type=int, default=65536)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-gpu-baseline", action="store_true")
    ap.add_argument("--no-fp32", action="store_true")
    ap.add_argument("--no-configs", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
