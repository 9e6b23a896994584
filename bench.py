"""Headline benchmark: articulatory frames/sec of the evaluation hot path (forward + P2CP + TVs).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --steps K --warmup W    # the reference's CPU path (oracle port)

Workload = BASELINE.json configs[1]: the encoder_decoder model at batch 256, sequence length 100,
followed by P2CP and the tract variables of prediction and target for every frame.  One "step" is
one such batch (25 600 frames).  Under torchrun every rank runs its own batch (utterance sharding,
weak scaling) and the fp64 metric accumulators are summed with one NCCL all-reduce per step.

Prints ONE JSON line (rank 0).  `value` = frames/s with inputs resident in HBM, timed with CUDA
events per step (L2 flushed between steps, outside the timed windows), max over ranks.  `e2e` =
the same step through the public API starting from pinned HOST buffers (H2D of tokens + targets and
D2H of the accumulators inside the timed region).
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


class ClockSampler:
    """SM clock + throttle reasons sampled every 2 ms through NVML while the timed region runs."""

    def __init__(self, index):
        self.index = index
        self.samples = []
        self.reasons = set()
        self.max_mhz = None
        self._stop = threading.Event()
        self.thread = None
        self.error = None

    def _run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            flags = {
                "hw_slowdown": nv.nvmlClocksEventReasonHwSlowdown if hasattr(nv, "nvmlClocksEventReasonHwSlowdown")
                else nv.nvmlClocksThrottleReasonHwSlowdown,
                "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown",
                                               getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0)),
                "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown",
                                               getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0)),
                "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap",
                                        getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0)),
            }
            get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons",
                                  getattr(nv, "nvmlDeviceGetCurrentClocksThrottleReasons", None))
            while not self._stop.is_set():
                self.samples.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                if get_reasons is not None:
                    r = int(get_reasons(h))
                    for name, bit in flags.items():
                        if bit and (r & int(bit)):
                            self.reasons.add(name)
                self._stop.wait(0.002)
        except Exception as exc:  # NVML missing: report it instead of inventing numbers
            self.error = repr(exc)

    def start(self):
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


def run_reference(args):
    """The reference's own CPU implementation of the path (oracle port: /root/reference does not
    exist on the GPU box), all host threads, on a bounded sample of the configs[1] workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import cpu_baseline as cb
    from oracle import model as omodel
    from artspeech_b200 import synthetic
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    sd = omodel.make_state_dict(VOCAB, N_ART, seed=0)
    sb, sl = args.ref_batch, args.seq_len
    tokens = synthetic.tokens(sb, sl, VOCAB, seed=0)
    lengths = [sl] * sb
    targets = synthetic.contours(sb * sl, N_ART, D, seed=1).view(sb, sl, N_ART, 2, D)
    for _ in range(max(args.warmup, 1) if args.warmup > 0 else 0):
        cb.evaluate_batch_cpu(sd, tokens[:2], lengths[:2], targets[:2])
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cb.evaluate_batch_cpu(sd, tokens, lengths, targets)
    dt = time.perf_counter() - t0
    value = args.steps * sb * sl / dt
    sample = f"{sb} utterances x {sl} frames per step (bounded sample of the 256 x {sl} batch), fp32, torch CPU ops"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"encoder_decoder BiGRU fwd + P2CP + TVs, batch 256 x len {sl} (configs[1])",
                   "sample": sample},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def run_ours(args):
    import torch.distributed as dist
    from artspeech_b200 import _lib, evaluation, models, synthetic

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

    B, T = args.batch, args.seq_len
    frames = B * T
    torch.manual_seed(0)                       # same initial weights as the reference constructor
    model = models.ArtSpeech(VOCAB, N_ART, precision=args.precision).to(dev).eval()
    # synthetic utterances of this rank (sharded by utterance: rank r owns its own batch)
    tokens_h = synthetic.tokens(B, T, VOCAB, seed=100 + rank).pin_memory()
    targets_h = synthetic.contours(frames, N_ART, D, seed=200 + rank).view(B, T, N_ART, 2, D).pin_memory()
    lengths = [T] * B
    tokens_d = tokens_h.to(dev)
    targets_d = targets_h.to(dev)
    uinc_d = synthetic.UPPER_INCISOR_REF.to(dev)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)  # > 126 MB L2

    stage_events = []

    def make_mark(store):
        def mark(name):
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            store.append((name, ev))
        return mark

    def step(tok, tgt, mark=None):
        res = evaluation.evaluate_batch(model, tok, lengths, tgt, upper_incisor=uinc_d, mark=mark)
        evaluation.reduce_accumulators(res["acc"])
        return res

    def barrier():
        if world > 1:
            dist.barrier()

    for _ in range(max(args.warmup, 3)):
        step(tokens_d, targets_d)
    torch.cuda.synchronize()
    l0 = _lib.launch_count()
    step(tokens_d, targets_d)
    launches_per_step = _lib.launch_count() - l0

    # The step is ~10 short kernels: replaying it from a CUDA graph removes the host launch gaps.
    graphed = None
    if not args.no_graph:
        graphed = evaluation.GraphedEvaluator(model, tokens_d, lengths, targets_d, upper_incisor=uinc_d)

    # The all-reduce of step i's accumulators is issued asynchronously on a copy of the vector, so
    # that it runs beside step i+1 (nothing needs the sum before the evaluation ends); the last timed
    # step waits for every outstanding collective before its end event is recorded.
    acc_ring = torch.zeros(max(args.steps, 1), N_ART + 14, dtype=torch.float64, device=dev)
    pending = []

    def fast_step(tok=None, tgt=None, slot=None):
        if graphed is None:
            return step(tokens_d if tok is None else tok, targets_d if tgt is None else tgt)
        res = graphed(tok, tgt)
        if slot is None or world == 1:
            evaluation.reduce_accumulators(res["acc"])
        else:
            acc_ring[slot].copy_(res["acc"])
            pending.append(dist.all_reduce(acc_ring[slot], op=dist.ReduceOp.SUM, async_op=True))
        return res

    for _ in range(max(args.warmup, 3)):
        fast_step()
    if world > 1:
        for k in range(2):
            fast_step(slot=k)
        for w in pending:
            w.wait()
        pending.clear()
    torch.cuda.synchronize()

    # ---------------- device-resident timing (value)
    sampler = ClockSampler(visible_index(local_rank))
    barrier(); torch.cuda.synchronize()
    sampler.start()
    starts, ends = [], []
    for i in range(args.steps):
        flush.zero_()                                   # evict L2 (not timed)
        s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
        s.record()
        fast_step(slot=i)
        if i == args.steps - 1:
            for w in pending:
                w.wait()                                # the main stream waits for every collective
            pending.clear()
        e.record()
        starts.append(s); ends.append(e)
    torch.cuda.synchronize(); barrier()
    step_ms = [s.elapsed_time(e) for s, e in zip(starts, ends)]
    total_ms = torch.tensor([sum(step_ms)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
    total_ms = float(total_ms.item())
    value = world * frames * args.steps / (total_ms * 1e-3)
    launches = launches_per_step * args.steps

    # ---------------- the same steps launched eagerly with CUDA events around every kernel
    model.profile(True)
    model.read_profile(reset=True)
    eager_ms = []
    for _ in range(args.steps):
        flush.zero_()
        s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
        marks = []
        s.record()
        step(tokens_d, targets_d, make_mark(marks))
        e.record()
        torch.cuda.synchronize()
        eager_ms.append(s.elapsed_time(e)); stage_events.append(marks)
    clocks = sampler.stop()
    model.profile(False)

    # per-stage device times (means over the timed steps)
    prof = model.read_profile(reset=True)
    stage_ms = {k: (v[0] / v[1] if v[1] else 0.0) for k, v in prof.items()}
    agg = {}
    for marks in stage_events:
        for (n0, e0), (n1, e1) in zip(marks[:-1], marks[1:]):
            agg[n1] = agg.get(n1, 0.0) + e0.elapsed_time(e1)
    for k in agg:
        agg[k] /= args.steps
    # The geometry stages of an eager step are host-launch-bound (a few short kernels behind Python
    # calls), so their kernels are timed on their own: L2 flush, event, call, event, queued ahead of
    # the device so that the interval holds the kernels only.
    from artspeech_b200.metrics import p2cp_distance
    from artspeech_b200.tract_variables import tract_variables_batch
    with torch.no_grad():
        outputs_d = model(tokens_d, lengths)
    ld = torch.tensor(lengths, dtype=torch.int32, device=dev)
    p2cp_d = p2cp_distance(outputs_d, targets_d)
    tvp_d = tract_variables_batch(outputs_d, synthetic.ARTICULATORS, uinc_d)[0]
    tvt_d = tract_variables_batch(targets_d, synthetic.ARTICULATORS, uinc_d)[0]

    def time_calls(fn):
        for _ in range(3):
            fn()
        pairs = []
        for _ in range(args.steps):
            flush.zero_()
            s_ = torch.cuda.Event(enable_timing=True); e_ = torch.cuda.Event(enable_timing=True)
            s_.record(); fn(); e_.record()
            pairs.append((s_, e_))
        torch.cuda.synchronize()
        return sum(a_.elapsed_time(b_) for a_, b_ in pairs) / len(pairs)

    stage_ms.update({
        "p2cp": time_calls(lambda: p2cp_distance(outputs_d, targets_d)),
        "tv_pred_and_target": time_calls(lambda: (tract_variables_batch(outputs_d, synthetic.ARTICULATORS, uinc_d),
                                                  tract_variables_batch(targets_d, synthetic.ARTICULATORS, uinc_d))),
        "accumulate": time_calls(lambda: evaluation.accumulate_device(p2cp_d, tvp_d, tvt_d, ld)),
        "eager_forward_total": agg.get("forward", 0.0), "eager_step": sum(eager_ms) / len(eager_ms)})

    # ---------------- roofline of the dominant kernel
    kernels = {
        "decoder": ("tensor", 2.0 * MAC_DECODER * frames, stage_ms.get("decoder", 0.0)),
        "gru_l0": ("tensor", 2.0 * MAC_RECURRENT / 2 * frames, stage_ms.get("gru_l0", 0.0)),
        "gru_l1": ("tensor", 2.0 * MAC_RECURRENT / 2 * frames, stage_ms.get("gru_l1", 0.0)),
        "proj_l1": ("tensor", 2.0 * MAC_PROJ_L1 * frames, stage_ms.get("proj_l1", 0.0)),
        "p2cp": ("hbm", float(BYTES_P2CP) * frames, stage_ms.get("p2cp", 0.0)),
        "tract_variables": ("hbm", float(BYTES_TV) * frames, stage_ms.get("tv_pred_and_target", 0.0)),
    }
    top = max(kernels, key=lambda k: kernels[k][2])
    bound, work, ms = kernels[top]
    if bound == "tensor":
        achieved = work / (ms * 1e-3) / 1e12 if ms > 0 else 0.0
        peak, unit = peaks["tflops_sustained"], "TFLOP/s"
    else:
        achieved = work / (ms * 1e-3) / 1e9 if ms > 0 else 0.0
        peak, unit = peaks["hbm_gbs"], "GB/s"
    # DRAM bytes per launch of each kernel from the committed `ncu --set full` capture of this
    # workload (profiles/ncu_traffic.json: dram__bytes_read.sum + dram__bytes_write.sum)
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if os.path.exists(tpath) and (B, T) == (256, 100):
        with open(tpath) as f:
            traffic = json.load(f).get("bytes_per_launch", {}).get(top)
    roofline = {"kernel": top, "bound": bound, "achieved": achieved, "peak": peak, "unit": unit,
                "frac": achieved / peak if peak else None,
                "frac_of_tf32_rate": (achieved / (0.5 * peak) if peak and bound == "tensor" else None),
                "traffic": traffic,
                "peak_source": f"{peaks['source']} (dense bf16 cuBLAS sustained; this kernel computes in tf32, whose tensor rate is half: frac_of_tf32_rate)"
                if bound == "tensor" else peaks["source"],
                "ms_per_launch": ms,
                "all_kernels_ms": {k: round(v[2], 4) for k, v in kernels.items()},
                "all_kernels_frac": {k: round((v[1] / (v[2] * 1e-3) / (1e12 if v[0] == "tensor" else 1e9)) /
                                              (peaks["tflops_sustained"] if v[0] == "tensor" else peaks["hbm_gbs"]), 4)
                                     if v[2] > 0 else None for k, v in kernels.items()}}

    # ---------------- area function on one step's worth of frames (configs[2] kernel; not part of `value`,
    # whose metric is fwd + P2CP/TV): synthetic walls, CUDA events, L2 flushed between launches
    from artspeech_b200 import area_function as af
    wi_d, we_d = (t.to(dev) for t in synthetic.walls(frames, seed=300 + rank))
    dgrid = af.DeviceGrid(af.build_semipolar_grid(**synthetic.AREA_GRID_ARGS), dev)
    for _ in range(3):
        af.area_function_batch(wi_d, we_d, dgrid)
    area_ms = []
    for _ in range(args.steps):
        flush.zero_()
        s_ = torch.cuda.Event(enable_timing=True); e_ = torch.cuda.Event(enable_timing=True)
        s_.record(); af.area_function_batch(wi_d, we_d, dgrid); e_.record()
        torch.cuda.synchronize()
        area_ms.append(s_.elapsed_time(e_))
    area_ms = sum(area_ms) / len(area_ms)
    area_gbs = BYTES_AREA * frames / (area_ms * 1e-3) / 1e9
    area_line = {"frames": frames, "ms": round(area_ms, 4), "frames_per_s": frames / (area_ms * 1e-3),
                 "bound": "hbm", "achieved_gbs": area_gbs, "frac": area_gbs / peaks["hbm_gbs"],
                 "bytes_per_frame": BYTES_AREA}

    # ---------------- end to end from pinned host buffers (e2e)
    # Every step copies its tokens + targets from pinned host memory and reads its accumulator
    # vector back; StreamedEvaluator overlaps the copy of step i+1 with the compute of step i.
    streamed = None if args.no_graph else evaluation.StreamedEvaluator(
        model, tokens_d, lengths, targets_d, upper_incisor=uinc_d, depth=2)

    def e2e_run(n):
        if streamed is not None:
            return streamed.run([(tokens_h, targets_h)] * n)
        accs = []
        for _ in range(n):
            r_ = fast_step(tokens_h, targets_h)         # H2D of tokens + targets from pinned memory
            accs.append(r_["acc"].cpu())                # D2H of the step's result (synchronises)
        return accs

    r = fast_step()
    torch.cuda.synchronize()
    e2e_run(3)
    barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    accs_h = e2e_run(args.steps)
    torch.cuda.synchronize()
    e2e_s = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    acc_h = accs_h[-1]
    assert len(accs_h) == args.steps and torch.equal(acc_h, r["acc"].cpu()), "streamed result differs"
    barrier()
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_value = world * frames * args.steps / float(e2e_s.item())
    h2d = tokens_h.numel() * tokens_h.element_size() + targets_h.numel() * targets_h.element_size()
    d2h = acc_h.numel() * acc_h.element_size()

    if old_affinity is not None:
        os.sched_setaffinity(0, old_affinity)

    # ---------------- CPU baseline (rank 0, single-GPU run only): oracle port on the host cores
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import cpu_baseline as cb
        cores = os.cpu_count() or 1
        sd = {k: v.detach().cpu() for k, v in model.state_dict().items()}
        v, secs = cb.time_cpu_baseline(sd, args.cpu_batch, T, repeats=1, threads=cores)
        cpu_baseline = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                        "sample": f"{args.cpu_batch} utterances x {T} frames of the same workload "
                                  f"({secs:.1f} s of host time), fp32 torch CPU ops, per-frame TV loop"}

    if rank == 0:
        summary = evaluation.summarize(r["acc"], N_ART)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None,
            "dtype": "tf32/bf16 mma, fp32 accumulate" if args.precision == "fast" else "f32 (3xTF32)",
            "data": "synthetic",
            "config": {"workload": f"encoder_decoder BiGRU fwd + P2CP + TVs (pred+target), batch {B} x len {T} "
                                   f"per GPU (configs[1]), 10 articulators x 50 points",
                       "precision": args.precision, "frames_per_step_per_gpu": frames,
                       "parallelism": f"utterance-sharded x{world}, one NCCL all-reduce of {N_ART + 14} fp64 per step "
                                      f"(asynchronous: it runs beside the next step)",
                       "l2": "256 MiB flush between timed steps; per-step working set > L2",
                       "launch": "CUDA graph replay of the captured step (target TVs on a forked branch beside the "
                                 "forward pass)" if graphed is not None else "eager"},
            "stage_ms": {k: round(v, 4) for k, v in stage_ms.items()},
            "roofline": roofline, "cpu_baseline": cpu_baseline,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "pipeline": "depth-2: H2D of step i+1 on a copy stream overlaps step i" if streamed is not None
                    else "sequential", "host_cpus_near_gpu": near_cpus},
            "area_function": area_line,
            "gpu_launches": int(launches), "clocks": clocks,
            "check": {"mean_p2cp_first_articulator": summary[synthetic.ARTICULATORS[0]]["p2cp"],
                      "frames_accumulated": summary["frames"]},
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--seq-len", type=int, default=100)
    ap.add_argument("--precision", default="fast", choices=["fast", "fp32"])
    ap.add_argument("--cpu-batch", type=int, default=32, help="utterances in the bounded CPU-baseline sample")
    ap.add_argument("--ref-batch", type=int, default=8, help="utterances per step of --impl reference")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="launch every step eagerly instead of replaying a CUDA graph")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
