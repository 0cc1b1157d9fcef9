"""kfb_fuse_noalloc: the gradient-gather entry point for callers that cannot
allocate device memory at the call site (plugins running inside another
framework's memory-planned step, latency-critical hooks, threads that may not
touch the allocator). It keeps its block partials in a static device array
inside the library instead of a caller buffer.

Called straight through the C ABI (ctypes on libkungfu_cuda.so) and compared
with the regular kfb_fuse on the same inputs, twice in a row (the array is
shared between calls and keeps the previous call's contents). This file is
collected last on purpose; it passes on its own and after the other tests."""
import ctypes as C
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'kungfu_b200', 'lib', 'libkungfu_cuda.so')
BF16, F32 = 1, 0
PTRS, I64S = C.POINTER(C.c_void_p), C.POINTER(C.c_int64)


def _lib():
    l = C.CDLL(LIB)
    l.kfb_last_error.restype = C.c_char_p
    l.kfb_fuse.restype = C.c_int
    l.kfb_fuse.argtypes = [PTRS, I64S, I64S, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    l.kfb_fuse_noalloc.restype = C.c_int
    l.kfb_fuse_noalloc.argtypes = [PTRS, I64S, I64S, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
    return l


def _case(seed, n_tensors, dtype):
    g = torch.Generator(device='cuda').manual_seed(seed)
    shapes = [(3 + (7 * i) % 50, 1 + (11 * i) % 37) for i in range(n_tensors)]
    ts = [(torch.randn(s, generator=g, device='cuda') * (1 + i % 3)).to(dtype) for i, s in enumerate(shapes)]
    offs, off = [], 0
    for t in ts:
        offs.append(off)
        off += (t.numel() + 15) // 16 * 16
    return ts, offs, off


def _arrays(ts, offs):
    p = (C.c_void_p * len(ts))(*[t.data_ptr() for t in ts])
    c = (C.c_int64 * len(ts))(*[t.numel() for t in ts])
    o = (C.c_int64 * len(ts))(*offs)
    return C.cast(p, PTRS), C.cast(c, I64S), C.cast(o, I64S), (p, c, o)


def test_fuse_noalloc_matches_fuse_two_calls_in_a_row():
    l = _lib()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    scratch = torch.zeros(1024, device='cuda')
    # second case is smaller than the first: stale partials of call 1 stay in the array and must not leak in
    for seed, n_tensors, dtype, code in ((1, 200, torch.bfloat16, BF16), (2, 37, torch.float32, F32)):
        ts, offs, total = _case(seed, n_tensors, dtype)
        pa, ca, oa, keep = _arrays(ts, offs)
        flat_ref = torch.zeros(total, dtype=dtype, device='cuda')
        flat_new = torch.zeros(total, dtype=dtype, device='cuda')
        ss_ref = torch.zeros(1, device='cuda')
        ss_new = torch.full((1, ), -1.0, device='cuda')
        rc = l.kfb_fuse(pa, ca, oa, n_tensors, code, flat_ref.data_ptr(), code, ss_ref.data_ptr(), scratch.data_ptr(), stream)
        assert rc == 0, l.kfb_last_error()
        rc = l.kfb_fuse_noalloc(pa, ca, oa, n_tensors, code, flat_new.data_ptr(), code, ss_new.data_ptr(), stream)
        assert rc == 0, l.kfb_last_error()
        torch.cuda.synchronize()
        assert torch.equal(flat_new, flat_ref)
        assert torch.allclose(ss_new, ss_ref, rtol=1e-4), (float(ss_new), float(ss_ref))
        oracle = sum((t.float() ** 2).sum() for t in ts)
        assert torch.allclose(ss_new[0], oracle, rtol=1e-4)


def test_fuse_noalloc_rejects_oversize():
    l = _lib()
    n = 1025  # KFB_FUSE_NOALLOC_MAX_TENSORS + 1
    p = (C.c_void_p * n)()
    c = (C.c_int64 * n)()
    o = (C.c_int64 * n)()
    out = torch.zeros(16, device='cuda')
    rc = l.kfb_fuse_noalloc(C.cast(p, PTRS), C.cast(c, I64S), C.cast(o, I64S), n, F32, out.data_ptr(), F32,
                            out.data_ptr(), C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc != 0
    assert b'exceed' in l.kfb_last_error()
