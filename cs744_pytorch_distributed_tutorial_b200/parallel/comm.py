"""Python face of the native rendezvous store, the symmetric heap and the NVLink collectives.

``Store``      -- TCP key-value store hosted by rank 0 at ``(master_ip, port)``; the role c10d's
                  TCPStore plays for the reference (ref: master/part2a/part2a.py:80-85).
``Communicator`` -- one cudaMalloc'd heap per rank, mapped into every peer with CUDA IPC, with a
                  bump allocator handing out torch tensors that sit at the *same byte offset* on
                  every rank.  Collectives address peers as ``peer_base[r] + offset`` from inside
                  the kernel (csrc/collectives.cu).
"""
from __future__ import annotations

import ctypes
from ctypes import POINTER, byref, c_char_p, c_double, c_float, c_int, c_longlong, c_void_p
from typing import List, Optional

import torch

from ..ops import _lib
from ..ops._lib import call

P, I, F, LL = c_void_p, c_int, c_float, c_longlong

_lib.declare("b200dp_store_create", [c_char_p, I, I, I, I, POINTER(c_void_p)])
_lib.declare("b200dp_store_set", [P, c_char_p, c_char_p, I])
_lib.declare("b200dp_store_get", [P, c_char_p, c_char_p, I, POINTER(c_int)])
_lib.declare("b200dp_store_add", [P, c_char_p, LL, POINTER(c_longlong)])
_lib.declare("b200dp_store_barrier", [P, c_char_p])
_lib.declare("b200dp_store_delete_prefix", [P, c_char_p])
_lib.declare("b200dp_store_destroy", [P])
_lib.declare("b200dp_comm_create", [P, I, I, I, LL, POINTER(c_void_p)])
_lib.declare("b200dp_comm_create_ex", [P, I, I, I, LL, I, POINTER(c_void_p)])
_lib.declare("b200dp_comm_has_multicast", [P])
_lib.declare("b200dp_comm_prefer_nvls", [P, I])
_lib.declare("b200dp_comm_destroy", [P])
_lib.declare("b200dp_comm_heap", [P, POINTER(c_void_p), POINTER(c_longlong), POINTER(c_longlong)])
_lib.declare("b200dp_comm_peer_base", [P, I, POINTER(c_void_p)])
_lib.declare("b200dp_comm_set_timeout", [P, c_double])
_lib.declare("b200dp_comm_barrier", [P, I, P])
_lib.declare("b200dp_allreduce", [P, LL, LL, LL, LL, LL, I, I, F, F, F, I, I, I, P])
_lib.declare("b200dp_ps_reduce_bcast", [P, LL, LL, LL, LL, I, I, I, I, P])
_lib.declare("b200dp_broadcast", [P, LL, LL, I, I, I, P])
_lib.declare("b200dp_allgather_slices", [P, LL, LL, LL, I, I, P])
_lib.declare("b200dp_reducer_create", [P, I, POINTER(c_longlong), POINTER(c_longlong), LL, LL, LL, I, I, F, F, F,
                                       I, I, POINTER(c_void_p)])
_lib.declare("b200dp_reducer_set_lr", [P, F])
_lib.declare("b200dp_reducer_stream", [P, POINTER(c_void_p)])
_lib.declare("b200dp_reducer_set_blocks", [P, I])
_lib.declare("b200dp_reducer_set_lazy_exit", [P, I])
_lib.declare("b200dp_reducer_set_staged", [P, LL, POINTER(c_longlong), POINTER(c_longlong), POINTER(c_longlong)])
_lib.declare("b200dp_reduce_staged_sgd", [P, LL, LL, LL, LL, LL, LL, LL, LL, LL, F, F, F, I, I, I, P])
_lib.declare("b200dp_reducer_bucket_ready", [P, I, P])
_lib.declare("b200dp_reducer_finalize", [P, P])
_lib.declare("b200dp_reducer_destroy", [P])

CH_MAIN = 0     # collectives issued on the compute stream
CH_BUCKET = 1   # per-bucket collectives on the reducer's side stream
CH_AUX = 2


class Store:
    """Client (and, on rank 0, host) of the rendezvous store."""

    def __init__(self, master_ip: str, port: int, rank: int, world: int, timeout_s: float = 120.0):
        self.rank, self.world = rank, world
        h = c_void_p()
        call("b200dp_store_create", master_ip.encode(), port, rank, world, int(timeout_s * 1000), byref(h))
        self._h = h

    @property
    def handle(self) -> c_void_p:
        return self._h

    def set(self, key: str, value: bytes) -> None:
        call("b200dp_store_set", self._h, key.encode(), value, len(value))

    def get(self, key: str, cap: int = 4096) -> bytes:
        buf = ctypes.create_string_buffer(cap)
        n = c_int(0)
        call("b200dp_store_get", self._h, key.encode(), buf, cap, byref(n))
        return buf.raw[: n.value]

    def add(self, key: str, delta: int = 1) -> int:
        out = c_longlong(0)
        call("b200dp_store_add", self._h, key.encode(), delta, byref(out))
        return out.value

    def barrier(self, name: str = "b") -> None:
        call("b200dp_store_barrier", self._h, name.encode())

    def delete_prefix(self, prefix: str) -> None:
        """Erase every key that starts with ``prefix`` (keys of exchanges every rank has finished)."""
        call("b200dp_store_delete_prefix", self._h, prefix.encode())

    def close(self) -> None:
        if self._h is None:
            return
        try:
            # rank 0 hosts the server: it must outlive every client's last request
            self.add("store/bye", 1)
            if self.rank == 0:
                import time
                deadline = time.time() + 10
                while self.add("store/bye", 0) < self.world and time.time() < deadline:
                    time.sleep(0.01)
        except Exception:
            pass
        call("b200dp_store_destroy", self._h)
        self._h = None


class _DeviceBlob:
    """Minimal ``__cuda_array_interface__`` carrier so torch can wrap heap memory without copying."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3,
                                         "strides": None}


class Communicator:
    """Symmetric heap + collectives for one rank.  ``world == 1`` works without a store."""

    def __init__(self, store: Optional[Store], rank: int, world: int, device: int, heap_bytes: int,
                 timeout_s: float = 20.0, multicast: bool = False):
        """``multicast=True`` allocates the heap through the CUDA virtual-memory API and maps an NVSwitch
        multicast view of it (NVLS all-reduce, ``algo=3``); the default heap is cudaMalloc + CUDA IPC."""
        self.rank, self.world, self.device = rank, world, device
        self.store = store
        h = c_void_p()
        call("b200dp_comm_create_ex", store.handle if store is not None else None, rank, world, device, heap_bytes,
             1 if multicast else 0, byref(h))
        self._h = h
        self.has_multicast = bool(_lib.lib().b200dp_comm_has_multicast(h))
        base, usable, data_off = c_void_p(), c_longlong(), c_longlong()
        call("b200dp_comm_heap", self._h, byref(base), byref(usable), byref(data_off))
        self.base_ptr = base.value
        self._cursor = data_off.value
        self._limit = data_off.value + usable.value
        self._blob = _DeviceBlob(self.base_ptr, self._limit)
        self._bytes = torch.as_tensor(self._blob, device=torch.device("cuda", device))
        call("b200dp_comm_set_timeout", self._h, float(timeout_s))
        self._keep: List[torch.Tensor] = []

    def prefer_nvls(self, on: bool = True) -> None:
        """Large all-reduces (and the bucket reducer's) use the in-switch multimem kernel."""
        call("b200dp_comm_prefer_nvls", self._h, int(on))

    # -- allocation --------------------------------------------------------------------
    def zeros(self, numel: int, dtype: torch.dtype = torch.float32, align: int = 256) -> torch.Tensor:
        """Allocate ``numel`` elements at the same offset on every rank (all ranks must call in the
        same order).  Heap memory is zero-initialised at creation."""
        itemsize = torch.empty((), dtype=dtype).element_size()
        off = (self._cursor + align - 1) // align * align
        nbytes = numel * itemsize
        if off + nbytes > self._limit:
            raise MemoryError(f"symmetric heap exhausted: need {nbytes} bytes at {off}, limit {self._limit}")
        self._cursor = off + nbytes
        t = self._bytes[off: off + nbytes].view(dtype)
        return t

    def offset_of(self, t: torch.Tensor) -> int:
        off = t.data_ptr() - self.base_ptr
        if not (0 <= off < self._limit):
            raise ValueError("tensor does not live in this communicator's symmetric heap")
        return off

    def peer_ptr(self, rank: int) -> int:
        b = c_void_p()
        call("b200dp_comm_peer_base", self._h, rank, byref(b))
        return b.value

    # -- collectives ---------------------------------------------------------------------
    def barrier(self, channel: int = CH_MAIN, stream: Optional[torch.cuda.Stream] = None) -> None:
        call("b200dp_comm_barrier", self._h, channel, _lib.stream_ptr(stream))

    def allreduce(self, grads: torch.Tensor, start: int, end: int, *, params: Optional[torch.Tensor] = None,
                  momentum_buf: Optional[torch.Tensor] = None, lr: float = 0.0, momentum: float = 0.0,
                  weight_decay: float = 0.0, prescale: bool = True, algo: int = 0, blocks: int = 0,
                  channel: int = CH_MAIN, stream: Optional[torch.cuda.Stream] = None) -> None:
        """Mean all-reduce of ``grads[start:end]`` (flat fp32 in the heap), optionally with the SGD
        update of ``params`` fused into the epilogue."""
        fuse = params is not None
        call("b200dp_allreduce", self._h, self.offset_of(grads), self.offset_of(params) if fuse else 0,
             self.offset_of(momentum_buf) if fuse else 0, start, end, int(fuse), int(prescale), lr, momentum,
             weight_decay, algo, blocks, channel, _lib.stream_ptr(stream))

    def ps_reduce_bcast(self, grads: torch.Tensor, start: int, end: int, *, root: int = 0,
                        staging: Optional[torch.Tensor] = None, blocks: int = 0, channel: int = CH_MAIN,
                        stream: Optional[torch.cuda.Stream] = None) -> None:
        push = staging is not None
        call("b200dp_ps_reduce_bcast", self._h, self.offset_of(grads), self.offset_of(staging) if push else 0, start,
             end, root, int(push), blocks, channel, _lib.stream_ptr(stream))

    def broadcast(self, t: torch.Tensor, root: int = 0, blocks: int = 0, channel: int = CH_MAIN,
                  stream: Optional[torch.cuda.Stream] = None) -> None:
        nbytes = t.numel() * t.element_size()
        if nbytes % 16:
            raise ValueError("broadcast needs a multiple of 16 bytes")
        call("b200dp_broadcast", self._h, self.offset_of(t), nbytes, root, blocks, channel, _lib.stream_ptr(stream))

    def allgather_slices(self, t: torch.Tensor, start: int, end: int, blocks: int = 0, channel: int = CH_MAIN,
                         stream: Optional[torch.cuda.Stream] = None) -> None:
        """Every rank stores the 1/W slice of ``t[start:end]`` it owns (the slicing of the fused two-shot /
        NVLS collectives) into every peer's copy -- assembles per-owner state such as SGD momentum."""
        call("b200dp_allgather_slices", self._h, self.offset_of(t), start, end, blocks, channel,
             _lib.stream_ptr(stream))

    def close(self) -> None:
        if self._h is not None:
            self._bytes = None
            call("b200dp_comm_destroy", self._h)
            self._h = None


class BucketReducer:
    """Side-stream per-bucket all-reduce(+SGD); see csrc/collectives.cu (Reducer)."""

    def __init__(self, comm: Communicator, buckets, grads: torch.Tensor, params: Optional[torch.Tensor],
                 momentum_buf: Optional[torch.Tensor], lr: float, momentum: float, weight_decay: float,
                 prescale: bool = True, blocks: int = 16, channel: int = CH_BUCKET):
        n = len(buckets)
        starts = (c_longlong * n)(*[b.start for b in buckets])
        ends = (c_longlong * n)(*[b.end for b in buckets])
        fuse = params is not None
        h = c_void_p()
        call("b200dp_reducer_create", comm._h, n, starts, ends, comm.offset_of(grads),
             comm.offset_of(params) if fuse else 0, comm.offset_of(momentum_buf) if fuse else 0, int(fuse),
             int(prescale), lr, momentum, weight_decay, blocks, channel, byref(h))
        self._h = h
        self.n_buckets = n

    def bucket_ready(self, k: int) -> None:
        call("b200dp_reducer_bucket_ready", self._h, k, _lib.stream_ptr())

    def finalize(self) -> None:
        call("b200dp_reducer_finalize", self._h, _lib.stream_ptr())

    def set_lr(self, lr: float) -> None:
        call("b200dp_reducer_set_lr", self._h, lr)

    def set_lazy_exit(self, on: bool = True) -> None:
        """On by default: only the last bucket kernel of a step keeps its exit barrier."""
        call("b200dp_reducer_set_lazy_exit", self._h, int(on))

    def set_blocks(self, blocks: int) -> None:
        """CTAs per bucket collective from the next ``bucket_ready`` on (re-capture the step afterwards)."""
        call("b200dp_reducer_set_blocks", self._h, int(blocks))

    def set_staged(self, stage_buf_off: int, pushed_lo, pushed_hi, stage_base) -> None:
        """Optional path: buckets whose conv weight gradient arrives through ``conv_wgrad_push`` staging run
        the owner-side ``reduce_staged_sgd`` kernel instead of the all-reduce."""
        n = self.n_buckets
        arr = lambda v: (c_longlong * n)(*[int(x) for x in v])
        call("b200dp_reducer_set_staged", self._h, int(stage_buf_off), arr(pushed_lo), arr(pushed_hi), arr(stage_base))

    def stream(self, device) -> torch.cuda.Stream:
        """The reducer's side stream as a torch stream: kernels queued on it after ``bucket_ready(k)``
        run behind bucket k's collective and are joined by ``finalize``."""
        s = c_void_p()
        call("b200dp_reducer_stream", self._h, byref(s))
        return torch.cuda.ExternalStream(s.value, device=device)

    def close(self) -> None:
        if self._h is not None:
            call("b200dp_reducer_destroy", self._h)
            self._h = None
