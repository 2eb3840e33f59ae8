"""Command line: the tutorial's three flags plus ``--mode`` and optional overrides.

    python -m cs744_pytorch_distributed_tutorial_b200.cli --mode part2b \
        --master-ip 10.10.1.1 --num-nodes 4 --rank 0

``--master-ip`` / ``--num-nodes`` / ``--rank`` keep the reference's spelling and meaning
(ref: master/part2a/part2a.py:136-143).  Unlike the reference they have defaults (127.0.0.1 / 1 / 0,
or torchrun's environment when present) and are validated.  One process per
node/GPU; launch W of them (by hand like the tutorial, or with torchrun).
"""
from __future__ import annotations

import argparse
import os
import sys

from .config import MODES, TrainConfig


def build_parser() -> argparse.ArgumentParser:
    env = os.environ
    p = argparse.ArgumentParser(prog="b200dp", description=__doc__, formatter_class=argparse.RawTextHelpFormatter)
    p.add_argument("--master-ip", dest="master_ip", type=str, default=env.get("MASTER_ADDR", "127.0.0.1"),
                   help="master ip, 10.10.1.1")
    p.add_argument("--num-nodes", dest="size", type=int, default=int(env.get("WORLD_SIZE", "1")),
                   help="number of nodes, 4")
    p.add_argument("--rank", dest="rank", type=int, default=int(env.get("RANK", "0")), help="rank, 0")
    p.add_argument("--mode", choices=MODES, default="part1")
    p.add_argument("--arch", default="VGG11")
    p.add_argument("--global-batch", type=int, default=256)
    p.add_argument("--per-rank-batch", type=int, default=None, help="weak scaling: fix the per-rank batch instead")
    p.add_argument("--iters", type=int, default=None, help="stop the epoch after this many iterations")
    p.add_argument("--epochs", type=int, default=1)
    p.add_argument("--lr", type=float, default=0.1)
    p.add_argument("--momentum", type=float, default=0.9)
    p.add_argument("--weight-decay", type=float, default=1e-4)
    p.add_argument("--seed", type=int, default=5000)
    p.add_argument("--port", type=int, default=None)
    p.add_argument("--device", choices=("auto", "cpu", "cuda"), default="auto")
    p.add_argument("--backend", choices=("auto", "gloo", "nccl", "native"), default="auto")
    p.add_argument("--bucket-plan", default="cap:12+last",
                   help="part3 gradient buckets: cap:<MiB> (greedy) | layer (one per conv layer) | ddp (torch's caps) | single")
    p.add_argument("--no-fuse-optimizer", action="store_true")
    p.add_argument("--nvls", choices=("auto", "on", "off"), default="auto",
                   help="NVSwitch multicast heap; in-switch all-reduce for large buckets (auto: on for world > 1)")
    p.add_argument("--sync-impl", choices=("auto", "oneshot", "twoshot", "nvls"), default="auto",
                   help="part2b all-reduce kernel (auto: by size and heap)")
    p.add_argument("--dtype", choices=("bf16", "tf32"), default="bf16",
                   help="native GPU path: bf16 operands / activations, or tf32 tensor-core math on fp32 activations")
    p.add_argument("--aggregate-eval", action="store_true",
                   help="also sum the per-rank correct counts of the test pass across ranks (reported by rank 0)")
    p.add_argument("--no-cuda-graph", action="store_true")
    p.add_argument("--train-size", type=int, default=50000)
    p.add_argument("--test-size", type=int, default=10000)
    p.add_argument("--real-data", action="store_true", help="train on ../../data/cifar-10-batches-py (an error if it is missing)")
    p.add_argument("--no-augment", action="store_true")
    p.add_argument("--loader", choices=("auto", "python", "native"), default="auto",
                   help="batch assembly: vectorised torch ops, or the C++ thread pool of csrc/loader.cpp")
    p.add_argument("--loader-threads", type=int, default=0, help="0 = auto (2 x CPUs / ranks, clamped to 4..8)")
    p.add_argument("--no-eval", action="store_true")
    p.add_argument("--checkpoint", type=str, default=None, help="save a checkpoint here after training")
    p.add_argument("--resume", type=str, default=None)
    return p


def config_from_args(argv=None) -> TrainConfig:
    a = build_parser().parse_args(argv)
    return TrainConfig(master_ip=a.master_ip, num_nodes=a.size, rank=a.rank, mode=a.mode, arch=a.arch,
                       global_batch=a.global_batch, per_rank_batch=a.per_rank_batch, iters=a.iters, epochs=a.epochs,
                       lr=a.lr, momentum=a.momentum, weight_decay=a.weight_decay, seed=a.seed, port=a.port,
                       device=a.device, backend=a.backend, bucket_plan=a.bucket_plan,
                       fuse_optimizer=not a.no_fuse_optimizer, nvls={"auto": None, "on": True, "off": False}[a.nvls], cuda_graph=not a.no_cuda_graph, sync_impl=a.sync_impl, dtype=a.dtype, aggregate_eval=a.aggregate_eval,
                       train_size=a.train_size, test_size=a.test_size, synthetic=not a.real_data,
                       augment=not a.no_augment, loader=a.loader, loader_threads=a.loader_threads, eval=not a.no_eval, checkpoint=a.checkpoint, resume=a.resume)


def main(argv=None) -> int:
    from .engine import Trainer
    cfg = config_from_args(argv)
    trainer = Trainer(cfg)
    try:
        trainer.fit()
    finally:
        trainer.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
