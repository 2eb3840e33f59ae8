"""Gradient synchronisation strategies on top of ``torch.distributed`` (gloo on CPU, nccl on
GPU).  These are the *plumbing / oracle* implementations: they reproduce the arithmetic of the
tutorial's four sync blocks for any world size and are what the gloo/CPU configuration and the
equivalence tests run.  The product path on B200 is ``sync_native`` (in-kernel NVLink
collectives); nothing here is on a timed GPU path.

The only thing that differs between the tutorial's parts is the block between ``backward()``
and ``optimizer.step()`` (layer L5) -- that block is a ``GradSync`` here.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch
import torch.distributed as dist


class GradSync:
    """Synchronise ``p.grad`` of ``params`` across the ranks of ``group``."""

    name = "none"
    syncs_buffers = False

    def __init__(self, params: Sequence[torch.nn.Parameter], world: int, rank: int, group=None):
        self.params = list(params)
        self.world, self.rank, self.group = world, rank, group

    def sync(self) -> None:  # pragma: no cover - interface
        raise NotImplementedError

    def pre_forward(self, module: torch.nn.Module) -> None:
        """Hook run before every forward (DDP broadcasts buffers here)."""

    def close(self) -> None:
        pass


class NoSync(GradSync):
    """part1: single process (ref: master/part1/part1.py:20-44)."""

    def sync(self) -> None:
        return


class GatherScatterSync(GradSync):
    """part2a: per parameter, gather every rank's gradient on rank 0, sum them in rank order
    starting from zero, divide by the world size and scatter the mean back in place
    (ref: master/part2a/part2a.py:42-52, slave/part2a/part2a.py:43-45)."""

    name = "gather_scatter"

    def sync(self) -> None:
        W = self.world
        for p in self.params:
            g = p.grad
            if self.rank == 0:
                parts = [torch.zeros_like(g) for _ in range(W)]
                dist.gather(g, parts, dst=0, group=self.group)
                total = torch.zeros_like(g)
                for r in range(W):
                    total += parts[r]
                mean = total / W
                dist.scatter(g, [mean for _ in range(W)], src=0, group=self.group)
            else:
                dist.gather(g, dst=0, group=self.group)
                dist.scatter(g, src=0, group=self.group)


class P2PParamServerSync(GradSync):
    """part2a_extra: the same parameter server built from isend/irecv, every transfer waited on
    in turn; rank 0 computes (g0 + g1 + ... ) / W (ref: master/part2a/part2a_extra.py:41-58,
    slave/part2a/part2a_extra.py:41-45)."""

    name = "p2p_param_server"

    def sync(self) -> None:
        W = self.world
        for p in self.params:
            g = p.grad
            if self.rank == 0:
                recvd = [torch.zeros_like(g) for _ in range(W - 1)]
                for r in range(1, W):
                    dist.irecv(recvd[r - 1], src=r).wait()
                total = g.clone()
                for t in recvd:
                    total = total + t
                mean = total / W
                g.copy_(mean)
                for r in range(1, W):
                    dist.isend(mean, dst=r).wait()
            else:
                dist.isend(g, dst=0).wait()
                dist.irecv(g, src=0).wait()


class AllReducePerParamSync(GradSync):
    """part2b: per parameter, pre-divide by the world size then all_reduce(SUM) in place
    (ref: master/part2b/part2b.py:43-45)."""

    name = "allreduce"

    def sync(self) -> None:
        for p in self.params:
            p.grad.div_(self.world)
            dist.all_reduce(p.grad, op=dist.ReduceOp.SUM, group=self.group)


class BucketedOverlapSync(GradSync):
    """part3: DDP-style bucketed all-reduce of pre-divided gradients, launched asynchronously as
    each bucket becomes ready during backward, waited on before the optimizer step; BatchNorm
    buffers are broadcast from rank 0 before every forward
    (ref: master/part3/part3.py:116; torch DDP semantics).

    ``buckets`` is a list of lists of parameters in gradient-ready order.  Parameters get a
    post-accumulate-grad hook; when all gradients of the next bucket in order have arrived the
    bucket is flattened, divided by W and all-reduced with ``async_op=True``.
    """

    name = "ddp_bucketed"
    syncs_buffers = True

    def __init__(self, params, world, rank, group=None, buckets: Optional[List[List[torch.nn.Parameter]]] = None):
        super().__init__(params, world, rank, group)
        self.buckets = buckets if buckets is not None else [list(reversed(self.params))]
        self._bucket_of = {}
        for bi, b in enumerate(self.buckets):
            for p in b:
                self._bucket_of[id(p)] = bi
        self._pending = [len(b) for b in self.buckets]
        self._ready = [False] * len(self.buckets)
        self._next = 0
        self._works = []
        self._hooks = [p.register_post_accumulate_grad_hook(self._on_grad) for p in self.params]
        self.launch_log: List[int] = []  # bucket launch order of the last iteration (for tests)

    def _on_grad(self, p: torch.nn.Parameter) -> None:
        bi = self._bucket_of[id(p)]
        self._pending[bi] -= 1
        if self._pending[bi] == 0:
            self._ready[bi] = True
            # buckets are launched strictly in order, like DDP's Reducer
            while self._next < len(self.buckets) and self._ready[self._next]:
                self._launch(self._next)
                self._next += 1

    def _launch(self, bi: int) -> None:
        ps = self.buckets[bi]
        flat = torch.cat([p.grad.reshape(-1) for p in ps]).div_(self.world)
        work = dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        self._works.append((work, flat, ps))
        self.launch_log.append(bi)

    def sync(self) -> None:
        """Finalize: wait for every bucket and copy the means back into ``p.grad``."""
        if self._next != len(self.buckets):
            # gradients that did not go through the hooks (e.g. set manually): launch the rest now
            for bi in range(self._next, len(self.buckets)):
                self._launch(bi)
        for work, flat, ps in self._works:
            work.wait()
            off = 0
            for p in ps:
                n = p.numel()
                p.grad.copy_(flat[off: off + n].view_as(p.grad))
                off += n
        self._works = []
        self._pending = [len(b) for b in self.buckets]
        self._ready = [False] * len(self.buckets)
        self._next = 0

    def begin_iteration(self) -> None:
        self.launch_log = []

    def pre_forward(self, module: torch.nn.Module) -> None:
        for b in module.buffers():
            dist.broadcast(b, src=0, group=self.group)

    def close(self) -> None:
        for h in self._hooks:
            h.remove()
        self._hooks = []


def broadcast_module_state(module: torch.nn.Module, group=None) -> None:
    """DDP constructor semantics: make every replica identical to rank 0."""
    for t in list(module.parameters()) + list(module.buffers()):
        dist.broadcast(t.data, src=0, group=group)


def all_reduce_scalar(value: float, group=None, device="cpu") -> float:
    """Sum a python scalar across ranks (the accuracy aggregation the reference intended with its
    unmatched ``isend(correct)``; ref: slave/part2b/part2b.py:68-69)."""
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return float(t.item())


def make_torch_sync(mode: str, params, world: int, rank: int, group=None, buckets=None) -> GradSync:
    if mode == "part1" or world == 1 and mode != "part3":
        return NoSync(params, world, rank, group)
    if mode == "part2a":
        return GatherScatterSync(params, world, rank, group)
    if mode == "part2a_extra":
        return P2PParamServerSync(params, world, rank, group)
    if mode == "part2b":
        return AllReducePerParamSync(params, world, rank, group)
    if mode == "part3":
        return BucketedOverlapSync(params, world, rank, group, buckets)
    raise ValueError(mode)
