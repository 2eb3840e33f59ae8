"""Training engine: ONE trainer for the tutorial's five scripts.

The reference copies the same loop into every script and varies only the block between
``backward()`` and ``optimizer.step()``.  Here that block is a pluggable strategy
and the rest is shared:

  * ``TorchBackend``  -- stock torch ops + ``torch.distributed`` (gloo on CPU, nccl on GPU).
    Runs without a GPU; it is the plumbing configuration and the numerics oracle.
  * ``NativeBackend`` -- the B200 path: sm_100a kernels for every layer, flat symmetric
    parameter/gradient buffers, in-kernel NVLink collectives with the SGD update fused into
    their epilogue, the whole step captured once in a CUDA graph and replayed.

``Trainer.train_epoch`` prints the reference's lines (``<batch> loss:  <value>`` every 20
iterations, ``average time:  <s>`` once; ref: master/part2a/part2a.py:54-59) and additionally
reports device-timed, max-over-ranks throughput as one JSON line.
"""
from __future__ import annotations

import os
import time
from typing import Dict, Iterable, List, Optional, Tuple

import numpy as np
import torch
import torch.nn as nn
import torch.optim as optim

from .config import TrainConfig
from .models.vgg import VGG
from .ops import native as K
from .utils.metrics import PhaseTimer, StepClock, emit_json, nvtx_range


def seed_everything(seed: int) -> None:
    """ref: master/part2a/part2a.py:89-90 -- same seed on every rank => identical initial weights."""
    torch.manual_seed(seed)
    np.random.seed(seed)


# =============================================================================================
# torch backend (CPU / gloo plumbing, NCCL comparison)
# =============================================================================================
class TorchBackend:
    def __init__(self, cfg: TrainConfig, model: VGG, device: torch.device):
        import torch.distributed as dist
        from .parallel import sync_torch
        self.cfg, self.model, self.device = cfg, model, device
        self.dist = dist
        self.owns_pg = False
        self.group = None
        if cfg.world_size > 1 or cfg.mode == "part3":
            if not dist.is_initialized():
                backend = cfg.backend if cfg.backend in ("gloo", "nccl") else ("nccl" if device.type == "cuda" else "gloo")
                os.environ["MASTER_ADDR"] = cfg.master_ip
                os.environ["MASTER_PORT"] = str(cfg.port)
                import datetime
                dist.init_process_group(backend, rank=cfg.rank, world_size=cfg.world_size,
                                        timeout=datetime.timedelta(seconds=cfg.timeout_s))
                self.owns_pg = True
            if cfg.mode in ("part2a", "part2b"):
                # the reference creates a sub-group holding every rank (ref: master/part2a/part2a.py:32)
                self.group = dist.new_group(list(range(cfg.world_size)))
        self.criterion = nn.CrossEntropyLoss().to(device)
        params = list(model.parameters())
        buckets = None
        if cfg.mode == "part3":
            sync_torch.broadcast_module_state(model, self.group)
            buckets = self._param_buckets(cfg.bucket_plan)
        self.sync = sync_torch.make_torch_sync(cfg.mode, params, cfg.world_size, cfg.rank, self.group, buckets)
        self.optimizer = optim.SGD(params, lr=cfg.lr, momentum=cfg.momentum, weight_decay=cfg.weight_decay)

    def _param_buckets(self, plan: str) -> List[List[nn.Parameter]]:
        from .models.native_vgg import FlatLayout
        named = list(self.model.named_parameters())
        lookup = dict(named)
        return [[lookup[n] for n in b.names] for b in FlatLayout(named).buckets(plan)]

    def train_step(self, data: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        data = data.to(self.device, non_blocking=True)
        target = target.to(self.device, non_blocking=True)
        self.model.train()
        self.sync.pre_forward(self.model)
        if hasattr(self.sync, "begin_iteration"):
            self.sync.begin_iteration()
        self.optimizer.zero_grad()
        out = self.model(data)
        loss = self.criterion(out, target)
        loss.backward()
        self.sync.sync()
        self.optimizer.step()
        return loss.detach()

    @torch.no_grad()
    def eval_batch(self, data: torch.Tensor, target: torch.Tensor) -> Tuple[torch.Tensor, int]:
        data, target = data.to(self.device), target.to(self.device)
        self.model.eval()
        self.sync.pre_forward(self.model)  # DDP also syncs buffers on eval forwards
        out = self.model(data)
        loss = self.criterion(out, target)
        correct = int((out.argmax(1) == target).sum().item())
        return loss, correct

    def input_event(self):
        """Event after which the host buffers passed to the last ``train_step`` may be overwritten."""
        if self.device.type != "cuda":
            return None
        ev = torch.cuda.Event()
        ev.record()
        return ev

    def all_reduce_sum(self, value: float) -> float:
        if self.cfg.world_size == 1 or not self.dist.is_initialized():
            return value
        from .parallel.sync_torch import all_reduce_scalar
        return all_reduce_scalar(value, None, self.device)

    def max_over_ranks(self, value: float) -> float:
        if self.cfg.world_size == 1 or not self.dist.is_initialized():
            return value
        t = torch.tensor([value], dtype=torch.float64, device=self.device)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def barrier(self) -> None:
        if self.cfg.world_size > 1 and self.dist.is_initialized():
            self.dist.barrier()

    def optimizer_state(self) -> dict:
        return self.optimizer.state_dict()

    def load_optimizer_state(self, sd: dict) -> None:
        self.optimizer.load_state_dict(sd)

    def synchronize(self) -> None:
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)

    def close(self) -> None:
        self.sync.close()
        if self.owns_pg and self.dist.is_initialized():
            self.dist.destroy_process_group()


# =============================================================================================
# native backend (B200)
# =============================================================================================
class NativeBackend:
    """Static-shape step on the sm_100a executor, captured in a CUDA graph.

    ``train_step(data, target)`` takes *host* (ideally pinned) tensors: the copy of the batch to
    the device is part of the step, issued on a copy stream one step ahead of the compute.
    """

    def __init__(self, cfg: TrainConfig, model: VGG, device: torch.device):
        from .parallel.comm import Communicator, Store
        from .parallel.sync_native import make_native_sync
        self.cfg, self.model, self.device = cfg, model, device
        torch.cuda.set_device(device)
        B = cfg.per_rank_batch
        self.store = None
        self.comm = None
        W = cfg.world_size
        need_comm = W > 1 or (cfg.mode == "part3" and cfg.single_rank_collectives)
        if W > 1:
            self.store = Store(cfg.master_ip, cfg.port, cfg.rank, W, timeout_s=cfg.timeout_s)
        heap = None
        if need_comm:
            n_params = sum(p.numel() for p in model.parameters())
            flat = (n_params + 4 * 64 + 1023) // 1024 * 1024 + 4096
            staging = (W - 1) * flat if cfg.mode == "part2a_extra" else 0
            if cfg.mode == "part3" and W > 1 and os.environ.get("B200DP_WGRAD_PUSH", "0") != "0":
                staging = flat + 64 * W * 16   # optional wgrad -> reduce-scatter staging slots
            heap_bytes = 4 * (3 * flat + staging) + (8 << 20)
            from .ops._lib import NativeError
            self.comm = None
            if cfg.use_nvls:
                # The multicast heap needs NVSwitch multicast + VMM support; the capability check fails
                # before any rank has talked to another, so every rank falls back together.
                try:
                    self.comm = Communicator(self.store, cfg.rank, W, device.index or 0, heap_bytes,
                                             timeout_s=cfg.timeout_s, multicast=True)
                    self.comm.prefer_nvls(True)
                except NativeError as e:
                    if cfg.nvls:      # explicitly requested: do not hide the failure
                        raise
                    if cfg.rank == 0:
                        print(f"b200dp: NVLS unavailable ({e}); using the CUDA-IPC heap and P2P collectives", flush=True)
                    self.comm = None
            if self.comm is None:
                self.comm = Communicator(self.store, cfg.rank, W, device.index or 0, heap_bytes, timeout_s=cfg.timeout_s)
            heap = self.comm
        self.native = model.enable_native(max_batch=B, heap=heap, dtype=cfg.dtype)
        # this backend always calls loss_fwd_bwd() after a training forward, so the classifier forward, the loss
        # and the classifier backward run as ONE launch (head_fwd_bwd) instead of three plus two memset nodes
        self.native.defer_head = True
        self.sync = make_native_sync(cfg.mode, self.native, self.comm, cfg)
        self.sync.on_init()
        # static step inputs / outputs
        self.x = torch.zeros(B, 3, 32, 32, device=device)
        Bp = (B + 1) // 2 * 2                    # label buffers padded to a multiple of 16 bytes (copy_pair)
        self._y_full = torch.zeros(Bp, dtype=torch.int64, device=device)
        self.y = self._y_full[:B]
        self._stage_x = [torch.zeros_like(self.x) for _ in range(2)]
        self._stage_y = [torch.zeros_like(self._y_full) for _ in range(2)]
        self._copy_stream = torch.cuda.Stream(device)
        self._copy_done = [torch.cuda.Event() for _ in range(2)]
        self._slot_free = [torch.cuda.Event() for _ in range(2)]
        self._slot = 0
        self._loss_host = torch.zeros(64, pin_memory=True)
        self._loss_events: List[Optional[torch.cuda.Event]] = [None] * 64
        self._loss_idx = 0
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self._last_input_event = None
        self.steps_done = 0
        self.launches_per_step = 0
        if cfg.cuda_graph:
            self._capture()

    # -- the step ------------------------------------------------------------------------------
    def _step_body(self) -> None:
        nv = self.native
        with nvtx_range("b200dp.forward"):
            self.sync.pre_forward()
            nv.forward(self.x, training=True)
            self.sync.post_forward()
        with nvtx_range("b200dp.loss"):
            nv.loss_fwd_bwd(self.y)
        with nvtx_range("b200dp.backward"):
            nv.backward()
        with nvtx_range("b200dp.sync_and_step"):
            self.sync.sync_and_step()
            if not self.sync.handles_shadows:
                nv.refresh_shadows()
            # within a captured step the auxiliary stream must be joined before the capture ends
            self.sync.pre_forward()

    def _capture(self) -> None:
        # warm up outside capture (module loading, lazy attribute setup), on a side stream as
        # torch requires; the warm-up steps train on zeros, so restore the state afterwards.
        nv = self.native
        keep = [t.clone() for t in (nv.flat_params, nv.flat_momentum, nv.bn_state)]
        s = torch.cuda.Stream(self.device)
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(2):
                self._step_body()
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize(self.device)
        from .ops import _lib
        self.graph = torch.cuda.CUDAGraph()
        n0 = _lib.launch_count
        with torch.cuda.graph(self.graph):
            self._step_body()
        self.launches_per_step = _lib.launch_count - n0
        torch.cuda.synchronize(self.device)
        for dst, src in zip((nv.flat_params, nv.flat_momentum, nv.bn_state), keep):
            dst.copy_(src)
        nv.refresh_shadows()
        torch.cuda.synchronize(self.device)
        if self.comm is not None and self.cfg.world_size > 1:
            self.store.barrier("post_capture")

    def stage_inputs(self, data: torch.Tensor, target: torch.Tensor) -> int:
        """Start the host->device copy of a batch on the copy stream; returns the staging slot."""
        slot = self._slot
        self._slot ^= 1
        with torch.cuda.stream(self._copy_stream):
            self._copy_stream.wait_event(self._slot_free[slot])
            self._stage_x[slot][: data.shape[0]].copy_(data, non_blocking=True)
            self._stage_y[slot][: target.shape[0]].copy_(target, non_blocking=True)
            self._copy_done[slot].record(self._copy_stream)
        self._last_input_event = self._copy_done[slot]
        return slot

    def run_staged(self, slot: int) -> torch.Tensor:
        self._eval_synced = False
        cur = torch.cuda.current_stream()
        cur.wait_event(self._copy_done[slot])
        K.copy_pair(self.x, self._stage_x[slot], self._y_full, self._stage_y[slot])
        self._slot_free[slot].record(cur)
        if self.graph is not None:
            self.graph.replay()
        else:
            self._step_body()
        self.steps_done += 1
        return self.native.loss

    def train_step(self, data: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        if data.shape[0] != self.cfg.per_rank_batch:
            return self._ragged_step(data, target)
        return self.run_staged(self.stage_inputs(data, target))

    def _ragged_step(self, data: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        """Partial last batch (the reference keeps it): same kernels, run eagerly."""
        nv = self.native
        x = data.to(self.device, non_blocking=True).float()
        y = target.to(self.device, non_blocking=True)
        self._last_input_event = torch.cuda.Event()
        self._last_input_event.record()
        self.sync.pre_forward()
        nv.forward(x, training=True)
        self.sync.post_forward()
        nv.loss_fwd_bwd(y)
        nv.backward()
        self.sync.sync_and_step()
        if not self.sync.handles_shadows:
            nv.refresh_shadows()
        self.sync.pre_forward()
        self.steps_done += 1
        return nv.loss

    def input_event(self):
        """Event after which the host buffers passed to the last ``train_step`` may be overwritten
        (recorded on the copy stream behind the host->device copy, or on the compute stream for a ragged batch)."""
        return self._last_input_event

    def read_loss_async(self) -> int:
        """Queue a device->host copy of the current loss; returns a ticket for ``loss_value``."""
        i = self._loss_idx
        self._loss_idx = (i + 1) % len(self._loss_events)
        self._loss_host[i: i + 1].copy_(self.native.loss.view(1), non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        self._loss_events[i] = ev
        return i

    def loss_value(self, ticket: int) -> float:
        self._loss_events[ticket].synchronize()
        return float(self._loss_host[ticket])

    @torch.no_grad()
    def eval_batch(self, data: torch.Tensor, target: torch.Tensor) -> Tuple[torch.Tensor, int]:
        nv = self.native
        x = data.to(self.device, non_blocking=True).float()
        y = target.to(self.device, non_blocking=True)
        if self.sync.syncs_buffers and not getattr(self, "_eval_synced", False):
            # DDP broadcasts buffers on eval forwards too; they cannot change during evaluation, so once
            self.sync.resync_buffers()
            self._eval_synced = True
        nv.forward(x, training=False)
        nv.loss_fwd_bwd(y, need_grad=False)
        return nv.loss.clone(), int(nv.correct.item())

    # -- small host-side reductions through the rendezvous store -------------------------------
    def _store_reduce(self, value: float, op: str) -> float:
        if self.store is None:
            return value
        import struct
        self._red_round = getattr(self, "_red_round", 0) + 1
        key = f"red/{self._red_round}/"
        self.store.set(key + str(self.cfg.rank), struct.pack("<d", value))
        vals = [struct.unpack("<d", self.store.get(key + str(r)))[0] for r in range(self.cfg.world_size)]
        if self.cfg.rank == 0 and self._red_round > 2:
            # every rank has read round - 2 (it has since written round - 1, which rank 0 just read)
            self.store.delete_prefix(f"red/{self._red_round - 2}/")
        return max(vals) if op == "max" else sum(vals)

    def all_reduce_sum(self, value: float) -> float:
        return self._store_reduce(value, "sum")

    def max_over_ranks(self, value: float) -> float:
        return self._store_reduce(value, "max")

    def barrier(self) -> None:
        torch.cuda.synchronize(self.device)
        if self.store is not None:
            self.store.barrier("engine")

    def optimizer_state(self) -> dict:
        """Collective at world > 1: every rank must call it (the momentum slices kept by their owners are
        gathered over NVLink first, so the dict equals what ``torch.optim.SGD`` would hold)."""
        nv = self.native
        self.sync.gather_optimizer_state()
        torch.cuda.synchronize(self.device)
        bufs = {e.name: nv._view(nv.flat_momentum, e).detach().cpu().contiguous() for e in nv.layout.entries}
        return {"momentum_buffers": bufs, "lr": self.cfg.lr, "momentum": self.cfg.momentum,
                "weight_decay": self.cfg.weight_decay}

    def load_optimizer_state(self, sd: dict) -> None:
        nv = self.native
        for e in nv.layout.entries:
            if e.name in sd.get("momentum_buffers", {}):
                nv._view(nv.flat_momentum, e).copy_(sd["momentum_buffers"][e.name])

    def synchronize(self) -> None:
        torch.cuda.synchronize(self.device)

    def close(self) -> None:
        torch.cuda.synchronize(self.device)
        self.graph = None
        self.sync.close()
        if self.comm is not None:
            self.comm.close()
        if self.store is not None:
            self.store.close()


# =============================================================================================
# Trainer
# =============================================================================================
class Trainer:
    def __init__(self, cfg: TrainConfig, model: Optional[VGG] = None):
        self.cfg = cfg
        dev = cfg.resolved_device()
        if dev == "cuda":
            local = int(os.environ.get("LOCAL_RANK", cfg.rank % max(torch.cuda.device_count(), 1)))
            self.device = torch.device("cuda", local)
            torch.cuda.set_device(self.device)
        else:
            self.device = torch.device("cpu")
            torch.set_num_threads(cfg.cpu_threads)  # ref: master/part2a/part2a.py:18
        seed_everything(cfg.seed)
        self.model = model if model is not None else VGG(cfg.arch)
        self.model.to(self.device)
        backend = cfg.backend
        if backend == "auto":
            backend = "native" if self.device.type == "cuda" else "gloo"
        self.backend_name = backend
        if backend == "native":
            from .models import native_vgg
            if self.device.type != "cuda" and not native_vgg._DRY_RUN:   # _DRY_RUN: CPU emulation tests only
                raise RuntimeError("the native backend needs a CUDA device")
            self.backend = NativeBackend(cfg, self.model, self.device)
        else:
            self.backend = TorchBackend(cfg, self.model, self.device)
        self.global_step = 0
        self.timer = PhaseTimer(self.device)
        if cfg.resume:
            self.load_checkpoint(cfg.resume)

    # -- training ------------------------------------------------------------------------------
    def train_step(self, data: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        loss = self.backend.train_step(data, target)
        self.global_step += 1
        return loss

    def train_epoch(self, loader: Iterable, epoch: int = 0) -> Dict[str, float]:
        """One pass over ``loader`` with the reference's console output plus device-side timing."""
        cfg = self.cfg
        clock = StepClock(self.device)
        ref_first = time.time()
        last_loss = float("nan")
        n_iters = 0
        for batch_idx, (data, target) in enumerate(loader):
            if cfg.iters is not None and batch_idx >= cfg.iters:
                break
            clock.step_begin(batch_idx)
            loss = self.train_step(data, target)
            if hasattr(loader, "consumed"):
                # the loader may refill this batch's pinned buffer once its host->device copy is done
                loader.consumed(self.backend.input_event())
            clock.step_end(batch_idx)
            n_iters += 1
            if batch_idx % cfg.log_interval == 0:
                last_loss = float(loss.item())
                print(batch_idx, "loss: ", last_loss, flush=True)
                if batch_idx == 0:
                    ref_first = time.time()
            if batch_idx == 10:
                # the reference divides a 10-iteration span by 9; we print the same
                # label but the unbiased value
                self.backend.synchronize()
                print("average time: ", (time.time() - ref_first) / 10, flush=True)
        self.backend.synchronize()
        stats = clock.summary(skip=min(3, max(n_iters - 1, 0)))
        stats["iters"] = n_iters
        stats["loss"] = last_loss
        if stats.get("s_per_iter"):
            stats["s_per_iter"] = self.backend.max_over_ranks(stats["s_per_iter"])
            stats["images_per_s"] = cfg.global_batch / stats["s_per_iter"]
        return stats

    @torch.no_grad()
    def evaluate(self, loader: Iterable) -> Dict[str, float]:
        """ref: master/part2a/part2a.py:62-77 -- mean of per-batch mean losses, top-1 accuracy, on the
        full (un-sharded) test set on every rank."""
        total_loss, correct, n_batches, n_samples = 0.0, 0, 0, 0
        for data, target in loader:
            loss, c = self.backend.eval_batch(data, target)
            total_loss += float(loss)      # host sync: the batch has been consumed
            correct += c
            n_batches += 1
            n_samples += int(target.shape[0])
        avg = total_loss / max(n_batches, 1)
        print('Test set: Average loss: {:.4f}, Accuracy: {}/{} ({:.0f}%)\n'.format(
            avg, correct, n_samples, 100. * correct / max(n_samples, 1)), flush=True)
        out = {"test_loss": avg, "correct": correct, "samples": n_samples, "accuracy": correct / max(n_samples, 1)}
        if self.cfg.aggregate_eval and self.cfg.world_size > 1:
            # what the reference's workers start and its master never finishes: the counts meet on rank 0
            # (ref: slave/part2b/part2b.py:68-69)
            total_c = int(round(self.backend.all_reduce_sum(float(correct))))
            total_n = int(round(self.backend.all_reduce_sum(float(n_samples))))
            if self.cfg.rank == 0:
                print("All ranks: Accuracy: {}/{} ({:.0f}%)\n".format(total_c, total_n, 100. * total_c / max(total_n, 1)),
                      flush=True)
            out.update({"correct_all_ranks": total_c, "samples_all_ranks": total_n})
        return out

    def fit(self) -> Dict[str, float]:
        """The tutorial's job: one epoch of training, then the test pass."""
        from .data import BatchPipeline, NativeBatchPipeline, build_dataset
        cfg = self.cfg
        train_set = build_dataset(cfg, True)
        test_set = build_dataset(cfg, False)
        # the C++ loader keeps up with the native step (380 k images/s end to end on one B200, bench.py
        # e2e_with_loader); the vectorised torch pipeline (~20 k images/s) is plenty for the CPU/gloo backend
        use_native = cfg.loader == "native" or (cfg.loader == "auto" and self.backend_name == "native")
        if use_native:
            train = NativeBatchPipeline(train_set, cfg.per_rank_batch, cfg.world_size, cfg.rank, train=True,
                                        augment=cfg.augment, drop_last=cfg.drop_last, seed=0,
                                        threads=cfg.loader_threads)
            test = NativeBatchPipeline(test_set, cfg.per_rank_batch, train=False, threads=cfg.loader_threads)
        else:
            train = BatchPipeline(train_set, cfg.per_rank_batch, cfg.world_size, cfg.rank, train=True,
                                  augment=cfg.augment, drop_last=cfg.drop_last, seed=0)
            test = BatchPipeline(test_set, cfg.per_rank_batch, train=False)
        out: Dict[str, float] = {}
        for epoch in range(cfg.epochs):
            train.sampler.set_epoch(epoch)
            out.update(self.train_epoch(train, epoch))
            if cfg.eval:
                out.update(self.evaluate(test))
        if cfg.checkpoint:
            self.save_checkpoint(cfg.checkpoint)
        if cfg.json_metrics and cfg.rank == 0:
            emit_json({"mode": cfg.mode, "backend": self.backend_name, "world_size": cfg.world_size,
                       "global_batch": cfg.global_batch, **out})
        return out

    # -- checkpoints -----------------------------------------------------------------------------
    def save_checkpoint(self, path: str) -> None:
        from .checkpoint import save_checkpoint
        self.backend.synchronize()
        opt = self.backend.optimizer_state()      # collective on the native backend (gathers sharded momentum)
        if self.cfg.rank == 0:
            save_checkpoint(path, self.model, opt, self.global_step)
        self.backend.barrier()

    def load_checkpoint(self, path: str) -> None:
        from .checkpoint import load_checkpoint
        ck = load_checkpoint(path, self.model)
        if ck.get("optimizer") is not None:
            self.backend.load_optimizer_state(ck["optimizer"])
        self.global_step = int(ck.get("step", 0))

    def close(self) -> None:
        self.backend.close()
