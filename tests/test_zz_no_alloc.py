"""The allocation-free BatchNorm+ReLU(+pool) backward, called through the C ABI.

``b200dp_bn_relu_pool_bwd_noalloc`` exists for hosts that may not allocate device memory at the
point where the backward runs (everything was handed to their own arena, a fixed memory plan, or
no tolerance for cudaMalloc's device-wide synchronisation), so they cannot pass the ``[2][C]``
``sums`` scratch that the regular reduce/apply pair needs.  The library keeps that scratch in a
built-in 16 KiB device array instead.  Here it is driven with raw pointers via ctypes, exactly as a
C caller would, and compared with the regular pair -- twice in a row on different data, because the
built-in array survives between calls and must be cleared by each of them.

This file sorts last on purpose: a run that stops at the first failure has then already executed
every other test.
"""
import ctypes
import gc
from ctypes import c_float, c_int, c_void_p

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _rel(a, b):
    return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-12)).item()


def _c_abi():
    from cs744_pytorch_distributed_tutorial_b200.ops import _lib
    lib = _lib.lib()
    P, I = c_void_p, c_int
    lib.b200dp_bn_relu_pool_bwd_noalloc.argtypes = [P, P, P, P, P, P, P, P, P, I, I, I, I, I, P]
    lib.b200dp_bn_relu_pool_bwd_noalloc.restype = c_int
    lib.b200dp_bn_relu_pool_bwd_reduce.argtypes = [P, P, P, P, P, P, P, I, I, I, I, I, P]
    lib.b200dp_bn_relu_pool_bwd_reduce.restype = c_int
    lib.b200dp_bn_relu_pool_bwd_apply.argtypes = [P, P, P, P, P, P, P, P, P, P, I, I, I, I, I, P]
    lib.b200dp_bn_relu_pool_bwd_apply.restype = c_int
    lib.b200dp_last_error.restype = ctypes.c_char_p
    return lib


def _case(seed, B, H, W, C, pool):
    g = torch.Generator(device=DEV).manual_seed(seed)
    y = (torch.randn(B, H, W, C, device=DEV, generator=g) * 1.3 + 0.2).bfloat16()
    Ho, Wo = (H // 2, W // 2) if pool else (H, W)
    dout = torch.randn(B, Ho, Wo, C, device=DEV, generator=g).bfloat16()
    y32 = y.float().view(-1, C)
    mean = y32.mean(0).contiguous()
    invstd = (y32.var(0, unbiased=False) + 1e-5).rsqrt().contiguous()
    gamma = (torch.rand(C, device=DEV, generator=g) + 0.5).contiguous()
    beta = (torch.randn(C, device=DEV, generator=g) * 0.2).contiguous()
    return y, dout, mean, invstd, gamma, beta


@pytest.mark.parametrize("shapes", [
    [(16, 8, 8, 256, False), (16, 8, 8, 256, True)],      # same C twice: stale sums would show
    [(8, 32, 32, 64, True), (32, 2, 2, 512, True)],       # different C between the two calls
])
def test_noalloc_matches_reduce_apply_on_two_calls_in_a_row(shapes):
    lib = _c_abi()
    stream = torch.cuda.current_stream().cuda_stream
    for call_idx, (B, H, W, C, pool) in enumerate(shapes):
        y, dout, mean, invstd, gamma, beta = _case(100 + call_idx, B, H, W, C, pool)
        # existing entry points: caller-owned, zeroed scratch
        sums = torch.zeros(2, C, device=DEV)
        dy_ref = torch.empty_like(y)
        dg_ref = torch.empty(C, device=DEV)
        db_ref = torch.empty(C, device=DEV)
        rc = lib.b200dp_bn_relu_pool_bwd_reduce(y.data_ptr(), dout.data_ptr(), mean.data_ptr(), invstd.data_ptr(),
                                                gamma.data_ptr(), beta.data_ptr(), sums.data_ptr(), B, H, W, C,
                                                int(pool), stream)
        assert rc == 0, lib.b200dp_last_error()
        rc = lib.b200dp_bn_relu_pool_bwd_apply(y.data_ptr(), dout.data_ptr(), mean.data_ptr(), invstd.data_ptr(),
                                               gamma.data_ptr(), beta.data_ptr(), sums.data_ptr(), dy_ref.data_ptr(),
                                               dg_ref.data_ptr(), db_ref.data_ptr(), B, H, W, C, int(pool), stream)
        assert rc == 0, lib.b200dp_last_error()
        # new entry point: no scratch argument, nothing allocated inside
        dy = torch.empty_like(y)
        dg = torch.empty(C, device=DEV)
        db = torch.empty(C, device=DEV)
        # "nothing allocated" is checked on the allocator's cumulative allocation COUNT: the byte total can move
        # on its own when Python collects objects of earlier tests (their frees would read as a change here)
        gc.collect()
        torch.cuda.synchronize()
        before = torch.cuda.memory_stats()["allocation.all.allocated"]
        rc = lib.b200dp_bn_relu_pool_bwd_noalloc(y.data_ptr(), dout.data_ptr(), mean.data_ptr(), invstd.data_ptr(),
                                                 gamma.data_ptr(), beta.data_ptr(), dy.data_ptr(), dg.data_ptr(),
                                                 db.data_ptr(), B, H, W, C, int(pool), stream)
        assert rc == 0, lib.b200dp_last_error()
        torch.cuda.synchronize()
        assert torch.cuda.memory_stats()["allocation.all.allocated"] == before
        # same tolerances as tests/test_gpu_kernels.py uses for this op
        assert _rel(dg, dg_ref) < 2e-3, f"call {call_idx}: dgamma {_rel(dg, dg_ref)}"
        assert _rel(db, db_ref) < 2e-3, f"call {call_idx}: dbeta {_rel(db, db_ref)}"
        assert _rel(dy, dy_ref) < 8e-3, f"call {call_idx}: dy {_rel(dy, dy_ref)}"


def test_noalloc_rejects_more_channels_than_the_builtin_scratch_holds():
    lib = _c_abi()
    C = 4096
    y = torch.zeros(1, 2, 2, C, device=DEV, dtype=torch.bfloat16)
    v = torch.ones(C, device=DEV)
    rc = lib.b200dp_bn_relu_pool_bwd_noalloc(y.data_ptr(), y.data_ptr(), v.data_ptr(), v.data_ptr(), v.data_ptr(),
                                             v.data_ptr(), y.data_ptr(), v.data_ptr(), v.data_ptr(), 1, 2, 2, C, 0,
                                             torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = lib.b200dp_last_error().decode()
    assert "2048" in msg and "4096" in msg
