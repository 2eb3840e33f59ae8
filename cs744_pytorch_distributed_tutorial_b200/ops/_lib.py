"""ctypes binding of ``libb200dp.so`` (the C ABI declared in ``include/b200dp.h``).

The library is loaded lazily.  On a machine with a GPU a missing or unloadable
library is a hard error -- there is no silent PyTorch fallback for the native
ops (the CPU/gloo plumbing path never calls in here).
"""
from __future__ import annotations

import ctypes
from ctypes import c_char_p, c_float, c_int, c_longlong, c_void_p
from typing import Optional

import torch

from . import _build

_lib: Optional[ctypes.CDLL] = None

P = c_void_p
I = c_int
F = c_float
LL = c_longlong

_PROTOTYPES = {
    "b200dp_conv_igemm": [P, P, P, P, P, I, I, I, I, I, I, I, P],
    "b200dp_conv_igemm_splitk": [P, P, P, P, P, P, I, I, I, I, I, I, I, I, P],
    "b200dp_conv_igemm_pair": [P, P, P, P, P, P, I, I, I, I, I, I, I, I, P],
    "b200dp_conv_dgrad_bn_reduce": [P, P, P, P, P, P, P, P, I, I, I, I, I, I, I, I, I, P],
    "b200dp_conv_igemm_ex": [P, P, P, P, P, P, I, I, I, I, I, I, I, I, I, I, P],
    "b200dp_conv_dgrad_bn_reduce_ex": [P, P, P, P, P, P, P, P, I, I, I, I, I, I, I, I, I, I, P],
    "b200dp_conv_wgrad_ex": [P, P, P, I, I, I, I, I, I, I, I, I, I, I, P],
    "b200dp_conv_wgrad": [P, P, P, I, I, I, I, I, I, I, I, I, P],
    "b200dp_conv_wgrad_tiled": [P, P, P, I, I, I, I, I, I, I, I, I, I, P],
    "b200dp_conv_wgrad_push": [P, P, P, I, I, I, I, I, I, I, I, P, I, I, I, P, LL, LL, LL, LL, P],
    "b200dp_im2col3x3_c3": [P, P, I, I, I, P],
    "b200dp_bn_stats": [P, P, LL, I, P],
    "b200dp_bn_relu_pool_fwd": [P, P, P, P, P, P, P, P, P, P, I, I, I, I, I, I, F, F, P],
    "b200dp_bn_relu_pool_bwd_reduce": [P, P, P, P, P, P, P, I, I, I, I, I, P],
    "b200dp_bn_relu_pool_bwd_apply": [P, P, P, P, P, P, P, P, P, P, I, I, I, I, I, P],
    "b200dp_bn_bwd_apply_nopool_tma": [P, P, P, P, P, P, P, P, P, P, I, I, I, I, P],
    "b200dp_bn_relu_pool_fwd_tma": [P, P, P, P, P, P, P, P, P, P, I, I, I, I, I, I, F, F, P],
    "b200dp_bn_bwd_reduce_from_out_tma": [P, P, P, P, P, LL, I, P],
    "b200dp_bn_bwd_apply_pool_tma": [P, P, P, P, P, P, P, P, P, P, I, I, I, I, P],
    "b200dp_bn_relu_pool_bwd_reduce_from_out": [P, P, P, P, P, LL, I, P],
    "b200dp_linear_fwd": [P, P, P, P, I, I, I, P],
    "b200dp_cross_entropy": [P, P, P, P, P, I, I, F, P],
    "b200dp_linear_bwd": [P, P, P, P, P, P, I, I, I, P],
    "b200dp_head_fwd_bwd": [P, P, P, P, P, P, P, P, P, P, P, I, I, I, F, P],
    "b200dp_sgd_momentum": [P, P, P, LL, F, F, F, F, P],
    "b200dp_im2col3x3_c3_ex": [P, P, I, I, I, I, P],
    "b200dp_im2col3x3_c3_ex2": [P, P, I, I, I, I, I, P],
    "b200dp_bn_stats_ex": [P, P, LL, I, I, P],
    "b200dp_bn_relu_pool_fwd_ex": [P, P, P, P, P, P, P, P, P, P, I, I, I, I, I, I, F, F, I, P],
    "b200dp_bn_relu_pool_bwd_reduce_ex": [P, P, P, P, P, P, P, I, I, I, I, I, I, P],
    "b200dp_bn_relu_pool_bwd_reduce_from_out_ex": [P, P, P, P, P, LL, I, I, P],
    "b200dp_bn_relu_pool_bwd_apply_ex": [P, P, P, P, P, P, P, P, P, P, I, I, I, I, I, I, P],
    "b200dp_linear_fwd_ex": [P, P, P, P, I, I, I, I, P],
    "b200dp_linear_bwd_ex": [P, P, P, P, P, P, I, I, I, I, P],
    "b200dp_head_fwd_bwd_ex": [P, P, P, P, P, P, P, P, P, P, P, I, I, I, F, I, P],
    "b200dp_fill_zero": [P, LL, P],
    "b200dp_copy_pair": [P, P, LL, P, P, LL, P],
    "b200dp_prep_conv_weights": [P, P, P, I, I, I, I, P],
    "b200dp_prep_conv_weights_batched": [I, P, P, P, P, P, P, P, P],
    "b200dp_prep_conv_weights_ex": [P, P, P, I, I, I, I, I, P],
    "b200dp_prep_conv_weights_batched_ex": [I, P, P, P, P, P, P, P, I, P],
}


class NativeError(RuntimeError):
    pass


# Entry points that do not launch a kernel (bookkeeping); everything else launches exactly one.
_NON_LAUNCHING = {"b200dp_store_create", "b200dp_store_set", "b200dp_store_get", "b200dp_store_add",
                  "b200dp_store_barrier", "b200dp_store_destroy", "b200dp_comm_create", "b200dp_comm_destroy",
                  "b200dp_comm_heap", "b200dp_comm_peer_base", "b200dp_comm_set_timeout", "b200dp_reducer_create",
                  "b200dp_comm_create_ex", "b200dp_comm_prefer_nvls", "b200dp_comm_has_multicast",
                  "b200dp_reducer_set_lr", "b200dp_reducer_finalize", "b200dp_reducer_destroy",
                  "b200dp_reducer_stream", "b200dp_reducer_set_blocks", "b200dp_reducer_set_staged",
                  "b200dp_reducer_set_lazy_exit"}
launch_count = 0  # kernels of this library launched (or captured) by this process so far


def lib() -> ctypes.CDLL:
    """Load (building first if the sources changed) and return the library."""
    global _lib
    if _lib is None:
        path = _build.build()
        handle = ctypes.CDLL(str(path), mode=ctypes.RTLD_GLOBAL)
        handle.b200dp_last_error.restype = c_char_p
        handle.b200dp_version.restype = c_char_p
        for name, argtypes in _PROTOTYPES.items():
            fn = getattr(handle, name, None)
            if fn is None:
                continue  # declared here before the translation unit exists; call() reports it
            fn.argtypes = argtypes
            fn.restype = c_int
        _lib = handle
    return _lib


def declare(name: str, argtypes: list) -> None:
    """Register the prototype of an entry point defined in another module's domain
    (collectives, runtime).  Safe to call before or after the library is loaded."""
    _PROTOTYPES[name] = argtypes
    if _lib is not None:
        fn = getattr(_lib, name)
        fn.argtypes = argtypes
        fn.restype = c_int


def available() -> bool:
    try:
        lib()
        return True
    except Exception:
        return False


def call(name: str, *args) -> None:
    """Invoke an ``int``-returning entry point; raise NativeError on non-zero."""
    fn = getattr(lib(), name, None)
    if fn is None:
        raise NativeError(f"{name} is not exported by libb200dp.so")
    rc = fn(*args)
    if name not in _NON_LAUNCHING:
        global launch_count
        launch_count += 1
    if rc != 0:
        msg = lib().b200dp_last_error().decode(errors="replace")
        raise NativeError(f"{name} failed (code {rc}): {msg}")


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    """Device pointer of a tensor (None -> NULL)."""
    if t is None:
        return None
    return t.data_ptr()


def stream_ptr(stream: Optional[torch.cuda.Stream] = None) -> int:
    s = stream if stream is not None else torch.cuda.current_stream()
    return s.cuda_stream


def version() -> str:
    return lib().b200dp_version().decode()
