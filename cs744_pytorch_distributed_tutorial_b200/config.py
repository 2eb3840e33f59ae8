"""Run configuration.

The reference has exactly three flags (``--master-ip``, ``--num-nodes``,
``--rank``; ref: master/part2a/part2a.py:137-141) and hard-codes everything else
as literals.  Every literal becomes a field here whose *default is the reference
value*, so running with only the three reference flags reproduces its setup
(SURVEY.md 5.6).
"""
from __future__ import annotations

import dataclasses
import os
from dataclasses import dataclass
from typing import Optional

MODES = ("part1", "part2a", "part2a_extra", "part2b", "part3")
SYNC_IMPLS = {"auto": 0, "oneshot": 1, "twoshot": 2, "nvls": 3}   # b200dp_allreduce algo codes

# MASTER_PORT literals (ref: master/part2a/part2a.py:83, master/part3/part3.py:72)
REFERENCE_PORTS = {"part1": 29501, "part2a": 29501, "part2a_extra": 29501, "part2b": 29501, "part3": 29508}


@dataclass
class TrainConfig:
    # --- the three reference flags -------------------------------------------------
    master_ip: str = "127.0.0.1"
    num_nodes: int = 1          # world size (one process per GPU / node)
    rank: int = 0
    # --- which of the tutorial's parts to run ---------------------------------------
    mode: str = "part1"
    # --- literals of the reference, now defaults -----------------------------------
    global_batch: int = 256     # ref: 256 in part1 (part1.py:17), 4 x 64 elsewhere (part2a.py:20)
    lr: float = 0.1             # ref: master/part1/part1.py:98
    momentum: float = 0.9
    weight_decay: float = 1e-4
    seed: int = 5000            # ref: master/part2a/part2a.py:89
    epochs: int = 1             # ref: master/part2a/part2a.py:130
    port: Optional[int] = None  # default: the reference port of the mode
    log_interval: int = 20      # ref: master/part2a/part2a.py:54
    num_workers: int = 2        # ref: master/part2a/part2a.py:109
    # part2b all-reduce kernel: auto (one-shot below 256 KiB, NVLS when the multicast heap exists, else
    # two-shot) | oneshot | twoshot | nvls
    sync_impl: str = "auto"
    # batch assembly for Trainer.fit: "python" = vectorised torch ops (data.BatchPipeline, ~20 k images/s),
    # "native" = C++ threads (csrc/loader.cpp, > 300 k images/s per thread); same sampling, different
    # augmentation draws
    loader: str = "auto"          # native next to a GPU (native backend), python with the torch/gloo backend
    loader_threads: int = 0     # 0 = auto: 2 * host CPUs / ranks, clamped to 4..8 (8 ranks on a 32-CPU box: 8 threads
                                # per rank fed 2.60 M images/s, 4 threads 1.87 M)
    cpu_threads: int = 4        # ref: master/part2a/part2a.py:18
    data_root: str = "../../data"
    # --- new knobs -------------------------------------------------------------------
    arch: str = "VGG11"
    device: str = "auto"        # auto | cpu | cuda
    # operands and stored activations of the native GPU path (fp32 master weights and gradients either way):
    # bf16 (tcgen05 kind::f16) | tf32 (tcgen05 kind::tf32 on fp32 activations -- the precision of the
    # reference's fp32 model with torch's TF32 convolutions)
    dtype: str = "bf16"
    backend: str = "auto"       # auto | gloo | nccl | native  (native = in-kernel NVLink collectives)
    # part3 gradient buckets: cap:<MiB> (greedy, closed at conv-layer boundaries) | layer (one per conv layer,
    # SURVEY 7.3) | ddp (torch's 1 MiB + 25 MiB caps) | single, each optionally with "+last" (the first conv block,
    # whose gradients come last, in a bucket of its own).  cap:12+last gives 18.9 / 14.2 / 3.9 MB + 8 KB on VGG-11:
    # each bucket kernel pays barrier round trips and a launch, so few large ones hide under backward where eight
    # did not, and the one collective that nothing can hide moves 8 KB instead of 4 MB
    bucket_plan: str = "cap:12+last"
    fuse_optimizer: bool = True  # SGD in the collective epilogue where the mode allows it
    # CTAs of a side-stream bucket kernel (part3); the GEMMs keep the other SMs.  Three 9.4 MB buckets
    # become ready in the first ~100 us of backward; 16 CTAs could not drain them before backward ended
    # (2 GPUs: 0.857 ms/step with 16, 0.771 with 32, 0.769 with 48 -- profiles/RESULTS.md section 0)
    comm_blocks: int = 32
    sync_blocks: int = 64       # CTAs of a collective issued after backward (part2a / part2b)
    # VMM heap + NVSwitch multicast; large all-reduces run in the switch (multimem).  None = automatic:
    # on whenever world_size > 1 (measured, part3 ms/step with NVLS vs the two-shot P2P kernel: 0.95 vs
    # 0.99 on 2 B200s, 0.90 vs 1.02 on 4 -- the in-switch reduction leaves the SMs and HBM to backward).
    nvls: Optional[bool] = None
    single_rank_collectives: bool = False  # world_size 1: still run the bucket reducer / comm kernels
    cuda_graph: bool = True
    iters: Optional[int] = None  # stop an epoch after this many iterations
    train_size: int = 50000     # CIFAR-10 sizes; synthetic data uses the same counts by default
    test_size: int = 10000
    synthetic: bool = True
    augment: bool = True
    drop_last: bool = False     # the reference keeps the partial last batch
    eval: bool = True
    checkpoint: Optional[str] = None
    resume: Optional[str] = None
    json_metrics: bool = True
    # every rank evaluates the full test set (as the reference does); with this flag the per-rank correct
    # counts are also summed across ranks and reported by rank 0 -- the intent of the workers' isend of
    # `correct` to rank 0 that the master never receives (ref: slave/part2b/part2b.py:68-69)
    aggregate_eval: bool = False
    per_rank_batch: Optional[int] = None    # set: weak scaling (global_batch follows); None: global_batch is split
    timeout_s: float = 120.0

    def __post_init__(self) -> None:
        if self.mode not in MODES:
            raise ValueError(f"mode must be one of {MODES}, got {self.mode!r}")
        if self.num_nodes < 1 or not (0 <= self.rank < self.num_nodes):
            raise ValueError(f"need 0 <= rank < num_nodes, got rank={self.rank} num_nodes={self.num_nodes}")
        if self.mode == "part1" and self.num_nodes != 1:
            raise ValueError("part1 is the single-process mode; use part2a/part2b/part3 for num_nodes > 1")
        if self.sync_impl not in SYNC_IMPLS:
            raise ValueError(f"sync_impl must be one of {tuple(SYNC_IMPLS)}, got {self.sync_impl!r}")
        if self.dtype not in ("bf16", "tf32"):
            raise ValueError(f"dtype must be 'bf16' or 'tf32', got {self.dtype!r}")
        if self.loader not in ("auto", "python", "native"):
            raise ValueError(f"loader must be 'auto', 'python' or 'native', got {self.loader!r}")
        if self.port is None:
            self.port = REFERENCE_PORTS[self.mode]
        if self.per_rank_batch is None:
            if self.global_batch % self.num_nodes != 0:
                raise ValueError(f"global batch {self.global_batch} is not divisible by {self.num_nodes} ranks")
            self.per_rank_batch = self.global_batch // self.num_nodes
        else:
            self.global_batch = self.per_rank_batch * self.num_nodes

    @property
    def world_size(self) -> int:
        return self.num_nodes

    @property
    def use_nvls(self) -> bool:
        return (self.num_nodes >= 2) if self.nvls is None else bool(self.nvls and self.num_nodes > 1)

    def resolved_device(self) -> str:
        if self.device != "auto":
            return self.device
        import torch
        return "cuda" if torch.cuda.is_available() else "cpu"

    def replace(self, **kw) -> "TrainConfig":
        d = dataclasses.asdict(self)
        d.update(kw)
        if "num_nodes" in kw or "global_batch" in kw:
            d.setdefault("per_rank_batch", None)
            if "per_rank_batch" not in kw:
                d["per_rank_batch"] = None
        return TrainConfig(**d)

    @staticmethod
    def from_env(**kw) -> "TrainConfig":
        """Fill rank / world size / master address from torchrun-style environment variables."""
        env = os.environ
        base = dict(
            master_ip=env.get("MASTER_ADDR", "127.0.0.1"),
            num_nodes=int(env.get("WORLD_SIZE", "1")),
            rank=int(env.get("RANK", "0")),
        )
        base.update(kw)
        return TrainConfig(**base)
