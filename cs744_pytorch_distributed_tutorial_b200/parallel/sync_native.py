"""Gradient synchronisation + optimizer strategies of the B200 path.

Each strategy is the block the tutorial puts between ``backward()`` and the end of
``optimizer.step()``, implemented with the in-kernel NVLink collectives of
``csrc/collectives.cu`` on the executor's flat symmetric buffers:

  part1         -> ``LocalSGD``: one fused SGD kernel (ref: master/part1/part1.py:38).
  part2a        -> ``ParamServerSync``: rank 0 pulls every rank's gradients over NVLink, sums in
                   rank order, divides by W and pushes the mean back; then the local SGD kernel
                   (ref: master/part2a/part2a.py:42-53).
  part2a_extra  -> the same with ``push=True``: workers store into rank 0's staging area first,
                   the isend/irecv data flow (ref: master/part2a/part2a_extra.py:41-58).
  part2b        -> ``AllReduceSync``: pre-divide, all-reduce, SGD -- by default ONE two-shot kernel
                   that reduce-scatters, applies SGD to the owned slice and all-gathers updated
                   parameters (ref: master/part2b/part2b.py:43-46).
  part3         -> ``BucketedOverlapSync``: the same fused kernel per bucket, launched on a side
                   stream the moment a bucket's last gradient is enqueued, overlapping the rest of
                   backward; BatchNorm buffers broadcast from rank 0 before every forward
                   (ref: master/part3/part3.py:116-117).

All calls are stream-ordered and host-sync free, hence capturable in the step's CUDA graph.
"""
from __future__ import annotations

from typing import Optional

import torch

from ..ops import native as K
from .comm import CH_AUX, CH_BUCKET, CH_MAIN, BucketReducer, Communicator


class NativeSync:
    name = "none"
    syncs_buffers = False
    handles_shadows = False   # True: the strategy refreshes the bf16 weight copies itself

    def __init__(self, native, comm: Optional[Communicator], lr: float, momentum: float, weight_decay: float,
                 fuse_optimizer: bool = True):
        self.native, self.comm = native, comm
        self.lr, self.momentum, self.weight_decay = lr, momentum, weight_decay
        self.fuse_optimizer = fuse_optimizer
        self.world = comm.world if comm is not None else 1
        self.n = native.layout.total

    def _local_sgd(self) -> None:
        nv = self.native
        K.sgd_momentum(nv.flat_params, nv.flat_grads, nv.flat_momentum, self.lr, self.momentum, self.weight_decay)

    def on_init(self) -> None:
        """Run once after construction, outside any graph capture."""

    def pre_forward(self) -> None:
        pass

    def post_forward(self) -> None:
        pass

    def resync_buffers(self) -> None:
        pass

    def sync_and_step(self) -> None:
        """Synchronise gradients and update parameters (everything after backward())."""
        raise NotImplementedError

    def sync_only(self) -> None:
        """Synchronise gradients into ``flat_grads`` without touching the parameters."""
        raise NotImplementedError

    def gather_optimizer_state(self) -> None:
        """Collective (every rank calls it): make ``flat_momentum`` complete on every rank.  The fused
        collectives update momentum only on the owner of each 1/W slice; strategies that shard it override this."""

    def close(self) -> None:
        pass


class LocalSGD(NativeSync):
    """One rank.  The optimizer runs in two phases: as soon as the gradients of the last ``split`` conv layers
    exist (95 % of all parameters live there and they are ready after about half of backward) one fused SGD
    kernel updates that range and one kernel refreshes their bf16 copies, in the shadow of the remaining
    backward; the small remainder follows after backward.  Phase 1 runs on its own stream, forked from the
    weight-gradient stream: queued on that stream it kept the weight gradients of the first three layers --
    the tail of the step -- waiting behind the optimizer traffic (in-graph timeline, profiles/round2).
    ``overlap=False``: one kernel after backward."""

    name = "none"

    def __init__(self, *a, overlap: bool = True, **kw):
        super().__init__(*a, **kw)
        self.overlap = bool(overlap)
        nv = self.native
        if self.overlap:
            self.buckets = nv.layout.buckets("layer")
            nv.set_bucket_plan(self.buckets)
            nv.on_bucket_ready = self._bucket_ready
            self.handles_shadows = True
            # phase 1 covers the buckets of the deepest blocks down to (and including) `first_block`
            n = len(nv.blocks)
            self.first_block = max(n - 5, 1)            # VGG-11: blocks 3..7 = 35.3 of 36.9 MB
            self.split_bucket = max(k for k in range(len(self.buckets)) if self.first_block in nv.blocks_of_bucket(k))
            self.split_end = self.buckets[self.split_bucket].end
            self._prep_hi = K.BatchedWeightPrep([b.prep for b in nv.blocks[self.first_block:]])
            self._prep_lo = K.BatchedWeightPrep([b.prep for b in nv.blocks[:self.first_block]])
            self._opt_stream = torch.cuda.Stream(nv.device)
            self._opt_fork, self._opt_join = torch.cuda.Event(), torch.cuda.Event()
            self._opt_pending = False

    def _sgd_range(self, lo: int, hi: int) -> None:
        nv = self.native
        K.sgd_momentum(nv.flat_params[lo:hi], nv.flat_grads[lo:hi], nv.flat_momentum[lo:hi], self.lr, self.momentum,
                       self.weight_decay)

    def _bucket_ready(self, k: int) -> None:
        if k != self.split_bucket:
            return
        # called on the weight-gradient stream right after wgrad(first_block): every gradient of phase 1 exists.
        # The data gradient of first_block is the last reader of the old bf16 copies of blocks >= first_block.
        nv = self.native
        self._opt_fork.record(torch.cuda.current_stream())
        self._opt_stream.wait_event(self._opt_fork)
        self._opt_stream.wait_event(nv._dgrad_done[self.first_block])
        with torch.cuda.stream(self._opt_stream):
            self._sgd_range(0, self.split_end)
            self._prep_hi()
            self._opt_join.record(self._opt_stream)
        self._opt_pending = True

    def sync_only(self) -> None:
        return

    def sync_and_step(self) -> None:
        if not self.overlap:
            self._local_sgd()
            return
        self._sgd_range(self.split_end, self.n)
        self._prep_lo()
        if self._opt_pending:
            torch.cuda.current_stream().wait_event(self._opt_join)
            self._opt_pending = False

    def close(self) -> None:
        if self.overlap:
            self.native.on_bucket_ready = None


class ParamServerSync(NativeSync):
    name = "gather_scatter"

    def __init__(self, *a, push: bool = False, blocks: int = 64, **kw):
        super().__init__(*a, **kw)
        self.push = push
        self.blocks = blocks
        self.staging = None
        if push:
            self.name = "p2p_param_server"
            self.staging = self.comm.zeros((self.world - 1) * self.n)

    def sync_only(self) -> None:
        self.comm.ps_reduce_bcast(self.native.flat_grads, 0, self.n, root=0, staging=self.staging, blocks=self.blocks,
                                  channel=CH_MAIN)

    def sync_and_step(self) -> None:
        self.sync_only()
        self._local_sgd()


class AllReduceSync(NativeSync):
    name = "allreduce"

    def __init__(self, *a, blocks: int = 64, algo: int = 0, **kw):
        super().__init__(*a, **kw)
        self.blocks, self.algo = blocks, algo

    def sync_only(self) -> None:
        self.comm.allreduce(self.native.flat_grads, 0, self.n, prescale=True, algo=self.algo, blocks=self.blocks,
                            channel=CH_MAIN)

    def sync_and_step(self) -> None:
        nv = self.native
        if self.fuse_optimizer:
            self.comm.allreduce(nv.flat_grads, 0, self.n, params=nv.flat_params, momentum_buf=nv.flat_momentum,
                                lr=self.lr, momentum=self.momentum, weight_decay=self.weight_decay, prescale=True,
                                algo=self.algo, blocks=self.blocks, channel=CH_MAIN)
        else:
            self.sync_only()
            self._local_sgd()

    def gather_optimizer_state(self) -> None:
        if self.fuse_optimizer and self.world > 1:
            self.comm.allgather_slices(self.native.flat_momentum, 0, self.n, channel=CH_MAIN)


class BucketedOverlapSync(NativeSync):
    name = "ddp_bucketed"
    syncs_buffers = True

    def __init__(self, *a, bucket_plan: str = "cap:12+last", blocks: int = 32, **kw):
        super().__init__(*a, **kw)
        nv = self.native
        self.buckets = nv.layout.buckets(bucket_plan)
        nv.set_bucket_plan(self.buckets)
        fuse = self.fuse_optimizer
        self.reducer = BucketReducer(self.comm, self.buckets, nv.flat_grads, nv.flat_params if fuse else None,
                                     nv.flat_momentum if fuse else None, self.lr, self.momentum, self.weight_decay,
                                     prescale=True, blocks=blocks, channel=CH_BUCKET)
        nv.on_bucket_ready = self.reducer.bucket_ready
        self.launch_log = []
        # Bucket kernels 0..n-2 of a step run without their exit barrier: what it guarantees is implied by the
        # entry barrier of the next bucket kernel on the channel; only the last one of a step keeps it
        # (measured on 2 B200s: 0.743 vs 0.773 ms/step).  The reducer does this by itself.  The operand copies
        # of the weights are therefore re-laid-out in one pass after the last bucket (the engine does it): doing it
        # per bucket behind each collective needs every bucket kernel's own exit barrier, which costs more at
        # 8 GPUs than the 18 us pass at the tail (0.664 vs 0.636 ms per step, profiles/round2/scale_8/ab.log).
        #
        # Optional (B200DP_WGRAD_PUSH=1): the weight-gradient GEMM of every 3x3 layer forwards its finished
        # tiles to the owning rank's staging slots as they complete (GEMM -> reduce-scatter in one kernel); the
        # bucket kernel on the reducer stream then only sums local memory, applies SGD and broadcasts the
        # parameters.
        import os
        self._wgrad_push = (os.environ.get("B200DP_WGRAD_PUSH", "0") != "0" and fuse and bucket_plan == "layer"
                            and self.world > 1 and self.world <= 8)
        if self._wgrad_push:
            self._setup_wgrad_push()

    def _setup_wgrad_push(self) -> None:
        import ctypes
        nv, W, comm = self.native, self.world, self.comm
        per4 = [(((b.end - b.start) // 4) + W - 1) // W for b in self.buckets]
        stage_base, total = [], 0
        for p4 in per4:
            stage_base.append(total)
            total += W * p4 * 4
        self._stage = comm.zeros(max(total, 4))
        stage_off = comm.offset_of(self._stage)
        self._peer_stage = (ctypes.c_void_p * W)(*[comm.peer_ptr(r) + stage_off for r in range(W)])
        pushed_lo = [b.start for b in self.buckets]
        pushed_hi = [b.start for b in self.buckets]
        nv.wgrad_push = {}
        self._push_counters = []
        for i, blk in enumerate(nv.blocks):
            if blk.first:
                continue          # Cin = 3: 27 columns, not 16-byte rows; stays on the all-reduce
            e = nv.layout.by_name[f"layers.{blk.spec.conv_idx}.weight"]
            k = next(j for j, b in enumerate(self.buckets) if b.start <= e.offset < b.end)
            bk = self.buckets[k]
            tiles = 9 * ((blk.spec.cout + 127) // 128) * (blk.spec.cin // 64)
            counters = torch.zeros(tiles, dtype=torch.int32, device=nv.device)
            self._push_counters.append(counters)
            nv.wgrad_push[i] = dict(counters=counters, world=W, rank=comm.rank, peer_stage=self._peer_stage,
                                    flat_off=e.offset, own_start=bk.start, own_len=bk.end - bk.start,
                                    stage_base=stage_base[k])
            pushed_lo[k], pushed_hi[k] = e.offset, e.offset + e.numel
        self.reducer.set_staged(stage_off, pushed_lo, pushed_hi, stage_base)

    def on_init(self) -> None:
        # DDP constructor: replicas start identical to rank 0
        nv = self.native
        if self.world > 1:
            self.comm.broadcast(nv.flat_params, root=0, channel=CH_MAIN)
            self.comm.broadcast(nv.bn_state, root=0, channel=CH_MAIN)
            nv.refresh_shadows()
            self._aux_stream = torch.cuda.Stream(nv.device)
            self._aux_fork = torch.cuda.Event()
            self._aux_join = torch.cuda.Event()
            self._aux_pending = False

    def pre_forward(self) -> None:
        """DDP ``_sync_buffers``: every forward starts from rank 0's BatchNorm state.  That state was
        final when rank 0 finished its *previous* forward, so the broadcast is issued there
        (``post_forward``) on an auxiliary stream and only joined here -- same values, no exposed wait."""
        if self.world > 1 and self._aux_pending:
            torch.cuda.current_stream().wait_event(self._aux_join)
            self._aux_pending = False

    def post_forward(self) -> None:
        if self.world > 1:
            nv = self.native
            cur = torch.cuda.current_stream()
            self._aux_fork.record(cur)
            self._aux_stream.wait_event(self._aux_fork)
            self.comm.broadcast(nv.bn_state, root=0, blocks=2, channel=CH_AUX, stream=self._aux_stream)
            self._aux_join.record(self._aux_stream)
            self._aux_pending = True

    def resync_buffers(self) -> None:
        """Synchronous form (evaluation after training, tests)."""
        if self.world > 1:
            self.pre_forward()
            self.comm.broadcast(self.native.bn_state, root=0, blocks=2, channel=CH_MAIN)

    def sync_only(self) -> None:
        self.reducer.finalize()

    def sync_and_step(self) -> None:
        self.reducer.finalize()
        if not self.fuse_optimizer:
            self._local_sgd()

    def gather_optimizer_state(self) -> None:
        # per bucket, with the bucket's own slicing; the small (one-shot) buckets already hold identical
        # momentum on every rank, for them the gather rewrites equal values
        if self.fuse_optimizer and self.world > 1:
            for b in self.buckets:
                self.comm.allgather_slices(self.native.flat_momentum, b.start, b.end, channel=CH_MAIN)

    def close(self) -> None:
        self.native.on_bucket_ready = None
        self.native.wgrad_push = None
        self.reducer.close()


def make_native_sync(mode: str, native, comm: Optional[Communicator], cfg) -> NativeSync:
    kw = dict(lr=cfg.lr, momentum=cfg.momentum, weight_decay=cfg.weight_decay, fuse_optimizer=cfg.fuse_optimizer)
    world = comm.world if comm is not None else 1
    # one rank: a mean over one gradient is the gradient -- no kernel needs to touch peer memory, so every
    # mode degenerates to the fused local SGD.  `single_rank_collectives` keeps the reducer / collective
    # path alive at world size 1 (tests, profiling of the comm kernels on one GPU).
    if mode == "part1" or (world == 1 and not (mode == "part3" and cfg.single_rank_collectives)):
        return LocalSGD(native, comm, **kw)
    if mode == "part2a":
        return ParamServerSync(native, comm, blocks=cfg.sync_blocks, **kw)
    if mode == "part2a_extra":
        return ParamServerSync(native, comm, push=True, blocks=cfg.sync_blocks, **kw)
    if mode == "part2b":
        from ..config import SYNC_IMPLS
        return AllReduceSync(native, comm, blocks=cfg.sync_blocks, algo=SYNC_IMPLS[getattr(cfg, "sync_impl", "auto")],
                             **kw)
    if mode == "part3":
        return BucketedOverlapSync(native, comm, bucket_plan=cfg.bucket_plan, blocks=cfg.comm_blocks, **kw)
    raise ValueError(mode)
