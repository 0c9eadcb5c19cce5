"""s2t_specaugment_static: SpecAugment's mean fill for hosts that cannot allocate device memory.

Some users of libs2t_b200.so are handed a feature batch on the GPU but are not allowed to call cudaMalloc (or torch's
allocator) for the small buffer of partial sums the mean fill needs: a data-loader worker or plugin that lives inside
another component's memory pool, runs while that component is capturing a stream, or works under a "no allocation after
start-up" rule.  For them the library carries the buffer itself, as a fixed array in its own image, and this entry point
uses it.  Nothing like it exists in the reference (its SpecAugment is numpy on the CPU).

The check calls the new entry point straight through the C ABI (ctypes, raw pointers) and compares the batch it produces
with the one the long-standing s2t_specaugment produces from a caller-supplied buffer: same kernels, same summation order,
so the suite's mean-fill bound of 1e-5 (tests/test_specaugment_gpu.py) is met with room to spare, and every element
outside the masks must be bit-identical.  Two calls are made back to back, the second on a smaller, different batch, so
that anything left in the shared array by the first call would show up if the second call read it.

This file sorts last on purpose: `pytest -x` stops at the first failure, and a failure here should not hide the others.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

MEAN_FILL_TOL = 1e-5          # the bound tests/test_specaugment_gpu.py uses for the mean fill


def _case(seed, lens, T, F, ld, n_f=2, n_t=2):
    g = torch.Generator().manual_seed(seed)
    B = len(lens)
    feats = torch.randn((B, T, ld), generator=g) * 3.0 + 1.5
    for b, n in enumerate(lens):
        feats[b, n:] = 0.0
    fm = torch.zeros((B, n_f, 2), dtype=torch.int32)
    tm = torch.zeros((B, n_t, 2), dtype=torch.int32)
    for b, n in enumerate(lens):
        for k in range(n_f):
            f = int(torch.randint(1, 28, (1,), generator=g))
            fm[b, k] = torch.tensor([int(torch.randint(0, F - f + 1, (1,), generator=g)), f])
        for k in range(n_t):
            t = int(torch.randint(1, max(2, min(100, n)), (1,), generator=g))
            tm[b, k] = torch.tensor([int(torch.randint(0, max(1, n - t + 1), (1,), generator=g)), t])
    return feats, torch.tensor(lens, dtype=torch.int32), fm, tm


def _run(lib, name, out, lens, fm, tm, F, use_mean, means=None):
    """One raw C-ABI call, in place on ``out``."""
    B, T, ld = out.shape
    args = [C.c_void_p(out.data_ptr()), 0, out.stride(1), out.stride(0), C.c_void_p(lens.data_ptr()), B, T, F,
            C.c_void_p(fm.data_ptr()), fm.shape[1], C.c_void_p(tm.data_ptr()), tm.shape[1], -0.75, int(use_mean)]
    if name == "s2t_specaugment":
        args.append(C.c_void_p(means.data_ptr()) if means is not None else None)
    args.append(C.c_void_p(torch.cuda.current_stream().cuda_stream))
    rc = getattr(lib, name)(*args)
    assert rc == 0, lib.s2t_last_error().decode()
    return out


def test_static_scratch_matches_caller_scratch_twice_in_a_row(cuda_lib):
    lib = cuda_lib
    F, ld = 80, 96
    cases = [_case(11, [400, 397, 250, 120, 33, 7], 400, F, ld),       # first call: 6 utterances
             _case(12, [150, 90, 1], 150, F, ld)]                      # second call: fewer, shorter, other data
    for use_mean in (1, 0):
        for feats, lens, fm, tm in cases:
            feats, lens, fm, tm = feats.cuda(), lens.cuda(), fm.cuda(), tm.cuda()
            means = torch.empty(feats.shape[0] * 32, device="cuda", dtype=torch.float32)
            want = _run(lib, "s2t_specaugment", feats.clone(), lens, fm, tm, F, use_mean, means)
            got = feats.clone()
            torch.cuda.synchronize()
            before = torch.cuda.memory_allocated()
            _run(lib, "s2t_specaugment_static", got, lens, fm, tm, F, use_mean)
            torch.cuda.synchronize()
            assert torch.cuda.memory_allocated() == before                     # torch's allocator was not involved
            changed = want != feats
            assert bool(changed.any())
            assert torch.equal(got[~changed], feats[~changed])                 # nothing outside the masks is touched
            assert float((got - want).abs().max()) <= (MEAN_FILL_TOL if use_mean else 0.0)
            assert torch.equal(got[..., F:], feats[..., F:])                   # columns beyond n_freqs untouched


def test_static_scratch_rejects_batches_beyond_its_size(cuda_lib):
    lib = cuda_lib
    too_many = 1025                                      # S2T_SPECAUGMENT_STATIC_MAX_BATCH + 1 (include/s2t_b200.h)
    feats = torch.zeros((1, 4, 80), device="cuda")
    lens = torch.zeros(too_many, dtype=torch.int32, device="cuda")
    fm = torch.zeros((too_many, 1, 2), dtype=torch.int32, device="cuda")
    rc = lib.s2t_specaugment_static(C.c_void_p(feats.data_ptr()), 0, 80, 320, C.c_void_p(lens.data_ptr()), too_many, 4, 80,
                                    C.c_void_p(fm.data_ptr()), 1, None, 0, 0.0, 1,
                                    C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == -1
    msg = lib.s2t_last_error().decode()
    assert "1025" in msg and "1024" in msg, msg
    torch.cuda.synchronize()                              # nothing was launched, nothing faulted
    assert float(feats.abs().max()) == 0.0
