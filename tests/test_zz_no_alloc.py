"""asb_eval_accumulate_static: the allocation-free variant of asb_eval_accumulate (built-in device scratch).

No counterpart in the reference.  Called through the C-ABI,
checked against the existing entry point on two calls in a row (the second with different inputs: the shared
scratch keeps the first call's contents and nothing of it may leak), plus the size limit.  Collected last on
purpose.
"""
import ctypes
import re
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

HEADER = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "artspeech_b200.h")


def _max_doubles():
    with open(HEADER) as f:
        return int(re.search(r"#define\s+ASB_EVAL_STATIC_MAX_DOUBLES\s+(\d+)", f.read()).group(1))


def _inputs(B, T, n_art, seed, dev):
    g = torch.Generator().manual_seed(seed)
    p2cp = torch.rand(B, T, n_art, generator=g).to(dev)
    tv_p = torch.rand(B, T, 4, generator=g).to(dev)
    tv_t = torch.rand(B, T, 4, generator=g).to(dev)
    lengths = torch.randint(1, T + 1, (B,), generator=g).sort(descending=True).values.to(torch.int32).to(dev)
    return p2cp, tv_p, tv_t, lengths


def _existing(lib, _lib, p2cp, tv_p, tv_t, lengths, n_art):
    B, T = p2cp.shape[:2]
    scratch = torch.empty(B * (n_art + 12), dtype=torch.float64, device=p2cp.device)
    acc = torch.full((n_art + 14,), float("nan"), dtype=torch.float64, device=p2cp.device)
    _lib.check(lib.asb_eval_accumulate(_lib.ptr(p2cp), _lib.ptr(tv_p), _lib.ptr(tv_t), _lib.ptr(lengths), B, T, n_art,
                                       _lib.ptr(scratch), _lib.ptr(acc), _lib.stream_ptr(p2cp.device)))
    return acc


def _static(lib, _lib, p2cp, tv_p, tv_t, lengths, n_art, acc):
    B, T = p2cp.shape[:2]
    return lib.asb_eval_accumulate_static(_lib.ptr(p2cp), _lib.ptr(tv_p), _lib.ptr(tv_t), _lib.ptr(lengths), B, T,
                                          n_art, _lib.ptr(acc), _lib.stream_ptr(p2cp.device))


def test_static_scratch_entry_equals_existing_entry_twice_in_a_row():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from artspeech_b200 import _lib
    lib = _lib.lib()
    dev = torch.device("cuda:0")
    n_art = 10
    # first call: a large batch, so that the second (smaller, different values) runs over stale scratch
    cases = [_inputs(256, 100, n_art, seed=11, dev=dev), _inputs(19, 33, n_art, seed=12, dev=dev)]
    wants = [_existing(lib, _lib, *c, n_art) for c in cases]
    # every buffer exists before the calls under test: they must not allocate, and torch's allocator is not
    # involved between them
    accs = [torch.full((n_art + 14,), float("nan"), dtype=torch.float64, device=dev) for _ in cases]
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated(dev)
    for c, acc in zip(cases, accs):
        _lib.check(_static(lib, _lib, *c, n_art, acc))
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated(dev) == before
    for want, acc in zip(wants, accs):
        # same kernels, same fixed summation order: the tolerance of test_accumulate_kernel_matches_torch, and in
        # fact the same bits
        assert torch.allclose(acc, want, rtol=1e-12, atol=1e-12)
        assert torch.equal(acc, want)
    # and against the plain torch aggregation the existing entry is tested with
    from artspeech_b200 import evaluation
    p2cp, tv_p, tv_t, lengths = cases[1]
    ref = evaluation.accumulate(p2cp, tv_p, tv_t, lengths.cpu().tolist())
    assert torch.allclose(accs[1], ref.to(dev), rtol=1e-12, atol=1e-12)


def test_static_scratch_entry_rejects_sizes_beyond_its_limit():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from artspeech_b200 import _lib
    lib = _lib.lib()
    dev = torch.device("cuda:0")
    n_art = 10
    limit = _max_doubles()
    B_ok = limit // (n_art + 12)
    acc = torch.zeros(n_art + 14, dtype=torch.float64, device=dev)
    big = _inputs(B_ok + 1, 2, n_art, seed=3, dev=dev)
    status = _static(lib, _lib, *big, n_art, acc)
    assert status == _lib.ASB_ERR_WORKSPACE
    msg = lib.asb_last_error_string().decode()
    assert "exceeds the built-in scratch" in msg and str(limit) in msg
    with pytest.raises(ValueError):
        _lib.check(status)
    # the largest accepted batch still agrees with the existing entry
    edge = _inputs(B_ok, 2, n_art, seed=4, dev=dev)
    _lib.check(_static(lib, _lib, *edge, n_art, acc))
    assert torch.equal(acc, _existing(lib, _lib, *edge, n_art))
    assert isinstance(status, int) and ctypes.sizeof(ctypes.c_double) * limit == 256 * 1024
