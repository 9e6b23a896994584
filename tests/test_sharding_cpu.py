"""Host-side logic of the multi-GPU path on CPU: utterance sharding + one all-reduce of the fp64
accumulators over gloo (world size 2) must equal the single-rank result."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from artspeech_b200 import evaluation, synthetic


def _per_frame_results(B, T, seed):
    g = torch.Generator().manual_seed(seed)
    p2cp = torch.rand(B, T, 10, generator=g)
    tv_p = torch.rand(B, T, 4, generator=g)
    tv_t = torch.rand(B, T, 4, generator=g)
    L = synthetic.sorted_lengths(B, 3, T, seed=seed)
    return p2cp, tv_p, tv_t, L


def _worker(rank, world, port, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    p2cp, tv_p, tv_t, L = _per_frame_results(12, 9, seed=3)
    mine = evaluation.shard_utterances(12, rank, world)
    acc = evaluation.accumulate(p2cp[mine], tv_p[mine], tv_t[mine], L[mine])
    evaluation.reduce_accumulators(acc)
    if rank == 0:
        torch.save(acc, out_path)
    dist.destroy_process_group()


def test_two_rank_reduction_equals_single_rank(tmp_path):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = str(tmp_path / "acc.pt")
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    got = torch.load(out)
    p2cp, tv_p, tv_t, L = _per_frame_results(12, 9, seed=3)
    want = evaluation.accumulate(p2cp, tv_p, tv_t, L)
    assert torch.allclose(got, want, rtol=1e-12, atol=1e-12)
    assert got[10].item() == 12 and got[23].item() == float(L.sum())


def test_accumulator_semantics():
    """run_test averages per-sentence means over sentences (evaluation.py:72-84,150-156)."""
    p2cp, tv_p, tv_t, L = _per_frame_results(5, 7, seed=1)
    acc = evaluation.accumulate(p2cp, tv_p, tv_t, L)
    want = sum(p2cp[b, :int(L[b])].double().mean(dim=0) for b in range(5))
    assert torch.allclose(acc[:10], want, rtol=1e-12)
    summary = evaluation.summarize(acc, 10)
    assert abs(summary[synthetic.ARTICULATORS[0]]["p2cp"] - want[0].item() / 5) < 1e-12
    assert summary["frames"] == float(L.sum())
    assert evaluation.shard_utterances(7, 1, 3) == [1, 4]
