#!/usr/bin/env python
"""Headline benchmark: VGG-11 / CIFAR-10-shaped training throughput (images/s) on N B200s.

    python bench.py --gpus 1 --steps 50 --warmup 5                       # this framework
    python bench.py --impl reference --gpus 1 --steps 50 --warmup 5      # the tutorial's own code
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29511 bench.py --gpus 8 --steps 50 --warmup 5

Metric and config are the ones BASELINE.json names: VGG-11, synthetic CIFAR-10-shaped data,
random-init weights, SGD(lr 0.1, momentum 0.9, wd 1e-4), mode part3 (bucketed all-reduce overlapped
with backward).  Default scaling is *weak* at 256 images per GPU, bf16 compute with fp32 master
weights (BASELINE.json config 5); ``--scaling strong`` fixes the global batch at 256 instead.

One JSON line is printed by rank 0.  ``value`` is the device-timed whole-job images/s over exactly
K steps (CUDA events, barrier + synchronize on both sides, max over ranks); ``e2e`` repeats the
measurement through ``Trainer.train_step`` with a pinned-host -> device copy of every step's batch
and a device -> host read of every step's loss inside the timed region.

Extra keys of the same line, same lease: ``strong`` -- the reference's own operating point, global batch 256
split over the N ranks (BASELINE.json configs 2-4), for part3 / part2b / part2a (one rank: part1), for BOTH
arms; ``same_precision`` -- the same step with ``--dtype tf32`` (fp32 activations, tf32 tensor-core math: the
precision of the reference arm); ``selfcheck`` (N > 1, this framework's arm) -- every collective kernel against a device-computed
expectation at VGG-11's bucket sizes plus 2,000 skewed launches, run before anything is timed (a mismatch
aborts with exit code 3); ``roofline`` -- the step's FLOPs and streamed bytes against MEASURED_PEAKS.json.
At one GPU there is no peer to synchronise with: both arms run the tutorial's single-process part1 loop.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import threading
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


# ------------------------------------------------------------------------------------------------
# clocks during the timed region (NVML; same fields as the nvidia-smi query of the profiling recipe)
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    REASONS = {
        0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
        0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
        0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting",
    }

    def __init__(self, index: int, period_s: float = 0.02):
        self.samples, self.reasons = [], set()
        self.max_mhz = None
        self._stop = threading.Event()
        self._thread = None
        self.period = period_s
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _sample(self):
        try:
            self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
            mask = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
            for bit, name in self.REASONS.items():
                if mask & bit and name != "gpu_idle":
                    self.reasons.add(name)
        except Exception:
            pass

    def start(self):
        if self.nv is None:
            return
        self._stop.clear()

        def loop():
            while not self._stop.is_set():
                self._sample()
                time.sleep(self.period)
        self._thread = threading.Thread(target=loop, daemon=True)
        self._thread.start()

    def stop(self):
        if self._thread is not None:
            self._sample()
            self._stop.set()
            self._thread.join()
            self._thread = None

    def summary(self) -> dict:
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}
        return {"sm_mhz": statistics.median(self.samples), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


def env_int(name: str, default: int) -> int:
    return int(os.environ.get(name, default))


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=50)
    p.add_argument("--warmup", type=int, default=5)
    p.add_argument("--impl", choices=("native", "reference"), default="native")
    p.add_argument("--mode", default="part3", choices=("part1", "part2a", "part2a_extra", "part2b", "part3"))
    p.add_argument("--scaling", choices=("weak", "strong"), default="weak")
    p.add_argument("--per-gpu-batch", type=int, default=256)
    p.add_argument("--global-batch", type=int, default=256, help="used with --scaling strong")
    p.add_argument("--bucket-plan", default="cap:12+last")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--comm-blocks", type=int, default=None)
    p.add_argument("--sync-blocks", type=int, default=None)
    p.add_argument("--port", type=int, default=None)
    p.add_argument("--nvls", choices=("auto", "on", "off"), default="auto",
                   help="multicast heap; large all-reduces reduce inside the NVSwitch (auto: on for world > 1)")
    p.add_argument("--single-rank-collectives", action="store_true",
                   help="at --gpus 1 still run the bucket reducer and comm kernels (default: fused local SGD)")
    p.add_argument("--dtype", choices=("bf16", "tf32"), default="bf16",
                   help="operand / activation precision of the native arm (fp32 master weights either way)")
    p.add_argument("--no-strong", action="store_true", help="skip the global-batch-256 runs of the other modes")
    p.add_argument("--no-tf32", action="store_true", help="skip the extra tf32 (same precision as the reference) run")
    p.add_argument("--no-selfcheck", action="store_true", help="skip the collective self-check at N > 1")
    p.add_argument("--no-loader", action="store_true",
                   help="skip the input-pipeline run (uint8 images -> augmentation -> training step)")
    p.add_argument("--loader-threads", type=int, default=0, help="0 = auto (2 x CPUs / ranks, clamped to 4..8)")
    return p.parse_args()


def working_set_mb(arch_blocks, batch: int, act_bytes: int = 2) -> float:
    """Bytes a training step streams through (activations fwd+bwd, weights/grads/momentum fp32)."""
    act = 0
    for s in arch_blocks:
        hw = s.hw * s.hw
        act += batch * hw * s.cout * act_bytes * 2                      # y and dy
        act += batch * (hw // 4 if s.pool else hw) * s.cout * act_bytes * 2  # out and dout
    params = 9_231_114 * 4 * 3 + 9_231_114 * act_bytes * 2           # fp32 p/g/m + operand copies Wf/Wd
    return (act + params) / 1e6


def step_gflop(arch_blocks, batch: int) -> float:
    """Forward + data-gradient + weight-gradient FLOPs of the conv and linear layers for one step."""
    f = 0.0
    for i, s in enumerate(arch_blocks):
        g = 2.0 * batch * s.hw * s.hw * s.cout * 9 * s.cin
        f += g * (2 if i == 0 else 3)                       # the first layer needs no data gradient
    f += 3 * 2.0 * batch * 512 * 10
    return f / 1e9


def roofline(arch_blocks, per_gpu_batch: int, ms_per_step: float, act_bytes: int) -> dict:
    """Achieved rate of one GPU's step against the driver-measured denominators (MEASURED_PEAKS.json)."""
    peaks = {}
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except Exception:
        pass
    source = "measured" if peaks else "fallback"
    tf = peaks.get("bf16_tflops_sustained", 1400.0)
    bw = peaks.get("hbm_gbs", 6650.0)
    gf = step_gflop(arch_blocks, per_gpu_batch)
    mb = working_set_mb(arch_blocks, per_gpu_batch, act_bytes)
    t = ms_per_step * 1e-3
    return {"step_gflop": round(gf, 1), "achieved_tflops": round(gf / t / 1e3, 1),
            "frac_of_bf16_sustained": round(gf / t / 1e3 / tf, 3), "streamed_mb": round(mb, 1),
            "achieved_gbs": round(mb / t / 1e3, 1), "frac_of_hbm": round(mb / t / 1e3 / bw, 3),
            "floor_ms": round(max(gf / tf, mb / bw), 4), "peaks": source}


def native_config(args, mode: str, world: int, rank: int, port: int, strong: bool):
    from cs744_pytorch_distributed_tutorial_b200.config import TrainConfig
    kw = dict(mode=mode, master_ip=os.environ.get("MASTER_ADDR", "127.0.0.1"), num_nodes=world, rank=rank,
              device="cuda", eval=False, bucket_plan=args.bucket_plan, port=port, dtype=args.dtype)
    kw["single_rank_collectives"] = args.single_rank_collectives
    kw["nvls"] = {"auto": None, "on": True, "off": False}[args.nvls]
    if args.comm_blocks:
        kw["comm_blocks"] = args.comm_blocks
    if args.sync_blocks:
        kw["sync_blocks"] = args.sync_blocks
    if strong:
        kw["global_batch"] = args.global_batch
    else:
        kw["per_rank_batch"] = args.per_gpu_batch
    return TrainConfig(**kw)


def time_replays(be, steps: int, sampler=None) -> float:
    """ms per step of `steps` replays of the captured step, device-timed, max over ranks."""
    be.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if sampler is not None:
        sampler.start()
    e0.record()
    for _ in range(steps):
        if be.graph is not None:
            be.graph.replay()
        else:
            be._step_body()
    e1.record()
    torch.cuda.synchronize()
    if sampler is not None:
        sampler.stop()
    be.barrier()
    return be.max_over_ranks(e0.elapsed_time(e1)) / steps


def make_batches(B: int, rank: int, n: int = 8):
    g = torch.Generator().manual_seed(1234 + rank)
    return [(torch.randn(B, 3, 32, 32, generator=g).pin_memory(),
             torch.randint(0, 10, (B,), generator=g).pin_memory()) for _ in range(n)]


def loader_e2e(tr, cfg, steps: int, threads: int, warm: int = 5) -> dict:
    """K training steps fed by ``NativeBatchPipeline`` from a uint8 dataset, augmentation on, the loss of every
    step read back.  Timed with CUDA events around the K steps (max over ranks)."""
    from cs744_pytorch_distributed_tutorial_b200.data import NativeBatchPipeline, SyntheticCIFAR10
    be = tr.backend
    B = cfg.per_rank_batch
    ds = SyntheticCIFAR10((warm + steps) * B * cfg.world_size, train=True)
    pipe = NativeBatchPipeline(ds, B, cfg.world_size, cfg.rank, train=True, augment=True, drop_last=True, seed=0,
                               threads=threads)
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    tickets, last, n = [], None, 0
    for i, (x, y) in enumerate(pipe):
        if i == warm:
            be.barrier()
            torch.cuda.synchronize()
            t0.record()
        tr.train_step(x, y)
        pipe.consumed(be.input_event())
        if i >= warm:
            n += 1
            tickets.append(be.read_loss_async())
            if len(tickets) > 8:
                be.loss_value(tickets.pop(0))
    last = [be.loss_value(t) for t in tickets][-1]
    t1.record()
    torch.cuda.synchronize()
    be.barrier()
    pipe.close()
    ms = be.max_over_ranks(t0.elapsed_time(t1)) / n
    return {"value": cfg.global_batch / (ms * 1e-3), "unit": "images/s", "ms_per_step": ms, "steps": n,
            "loader": f"csrc/loader.cpp, {pipe.threads} threads per rank, RandomCrop(32,4)+flip+normalise, pinned ring",
            "host_cpus": os.cpu_count(),
            "last_loss": last}


def run_native(args) -> dict:
    from cs744_pytorch_distributed_tutorial_b200.config import TrainConfig
    from cs744_pytorch_distributed_tutorial_b200.engine import Trainer
    from cs744_pytorch_distributed_tutorial_b200.ops import _lib

    world = env_int("WORLD_SIZE", 1)
    rank = env_int("RANK", 0)
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}; launch with torch.distributed.run for N>1")
    base_port = args.port or env_int("MASTER_PORT", 29500) + 37
    torch.cuda.set_device(env_int("LOCAL_RANK", rank))

    # ---- N > 1: prove the collective kernels on this box before anything is timed
    selfcheck = None
    if world > 1 and not args.no_selfcheck:
        from cs744_pytorch_distributed_tutorial_b200.parallel.comm import Store
        from cs744_pytorch_distributed_tutorial_b200.parallel.selfcheck import run_selfcheck
        store = Store(os.environ.get("MASTER_ADDR", "127.0.0.1"), base_port + 20, rank, world, timeout_s=120)
        res = run_selfcheck(store, rank, world, torch.cuda.current_device(), multicast=args.nvls != "off")
        bad = store.add("selfcheck/failed", 0 if res["ok"] else 1)
        store.barrier("selfcheck")
        bad = store.add("selfcheck/failed", 0)
        store.close()
        selfcheck = {"ok": bad == 0, "cases": res["cases"], "seconds": res["seconds"], "nvls": res["nvls"]}
        if bad:
            if not res["ok"]:
                print("\n".join(res["failures"][:10]), file=sys.stderr, flush=True)
            if rank == 0:
                print(json.dumps({"impl": "native", "selfcheck": selfcheck, "error": "collective self-check failed"}),
                      flush=True)
            raise SystemExit(3)

    # one GPU has no peer: every mode is the tutorial's single-process loop (part1), and is labelled so
    mode = args.mode if (world > 1 or args.single_rank_collectives) else "part1"
    strong_main = args.scaling == "strong"
    cfg = native_config(args, mode, world, rank, base_port, strong_main)
    tr = Trainer(cfg)
    be = tr.backend
    B = cfg.per_rank_batch
    host_batches = make_batches(B, rank)
    sampler = ClockSampler(torch.cuda.current_device())
    act_bytes = 4 if args.dtype == "tf32" else 2

    # ---- device-timed: K replays of the captured step, inputs already resident on the device
    for i in range(max(args.warmup, 3)):
        be.run_staged(be.stage_inputs(*host_batches[i % 8]))
    n0 = _lib.launch_count
    ms_dev = time_replays(be, args.steps, sampler)
    launches = be.launches_per_step * args.steps if be.graph is not None else _lib.launch_count - n0

    # ---- end to end through the public API: H2D of every batch + D2H of every loss in the timed region
    e2e = None
    if not args.no_e2e:
        for i in range(3):
            tr.train_step(*host_batches[i % 8])
        be.barrier()
        torch.cuda.synchronize()
        t0 = torch.cuda.Event(enable_timing=True)
        t1 = torch.cuda.Event(enable_timing=True)
        t0.record()
        tickets = []
        for i in range(args.steps):
            tr.train_step(*host_batches[i % 8])
            tickets.append(be.read_loss_async())
            if len(tickets) > 8:                      # keep the host at most a few steps ahead
                be.loss_value(tickets.pop(0))
        last = [be.loss_value(t) for t in tickets][-1]
        t1.record()
        torch.cuda.synchronize()
        be.barrier()
        ms_e2e = be.max_over_ranks(t0.elapsed_time(t1)) / args.steps
        e2e = {"value": cfg.global_batch / (ms_e2e * 1e-3), "unit": "images/s", "ms_per_step": ms_e2e,
               "h2d_bytes_per_step": B * (3 * 32 * 32 * 4 + 8), "d2h_bytes_per_step": 4, "last_loss": last}
    # ---- end to end from uint8 images: the C++ loader (sampling, crop, flip, normalise into pinned ring buffers)
    # ---- feeds Trainer.train_step; the counterpart of the reference's DataLoader run
    e2e_loader = None
    if not args.no_loader:
        try:
            e2e_loader = loader_e2e(tr, cfg, min(args.steps, 200), args.loader_threads)
        except Exception as e:
            e2e_loader = {"unavailable": f"{type(e).__name__}: {e}"[:200]}
    specs = tr.model.block_specs()
    out = {
        "metric": "VGG-11 CIFAR-10 images/sec", "impl": "native", "value": cfg.global_batch / (ms_dev * 1e-3),
        "unit": "images/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_dev, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
        "dtype": args.dtype, "data": "synthetic (random CIFAR-10-shaped tensors, random-init weights)",
        "config": {"model": "VGG-11", "global_batch": cfg.global_batch, "per_gpu_batch": B, "seq_len": None,
                   "image": "3x32x32", "mode": mode, "parallelism": f"dp{world}", "optimizer": "SGD lr0.1 m0.9 wd1e-4",
                   "bucket_plan": args.bucket_plan, "cuda_graph": be.graph is not None, "nvls": cfg.use_nvls,
                   "l2": f"no explicit flush: one step streams ~{working_set_mb(specs, B, act_bytes):.0f} MB "
                         f"(> 126 MB L2) and rewrites every activation and parameter"},
        "gpu_launches": launches, "clocks": sampler.summary(),
        "roofline": roofline(specs, B, ms_dev, act_bytes),
    }
    if e2e is not None:
        out["e2e"] = e2e
    if e2e_loader is not None:
        out["e2e_with_loader"] = e2e_loader
    if selfcheck is not None:
        out["selfcheck"] = selfcheck
    if world > 1:
        # exposed communication = step time - the same step with the sync block replaced by the local
        # fused SGD, measured on every rank with an identical no-communication trainer
        local = Trainer(TrainConfig(mode="part1", num_nodes=1, rank=0, device="cuda", eval=False, per_rank_batch=B,
                                    dtype=args.dtype))
        lb = local.backend
        for i in range(3):
            lb.run_staged(lb.stage_inputs(*host_batches[i % 8]))
        be.barrier()
        torch.cuda.synchronize()
        l0, l1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0.record()
        for _ in range(args.steps):
            lb.graph.replay()
        l1.record()
        torch.cuda.synchronize()
        ms_local = be.max_over_ranks(l0.elapsed_time(l1)) / args.steps
        local.close()
        out["compute_only_ms_per_step"] = ms_local
        out["exposed_comm_ms_per_step"] = ms_dev - ms_local
        out["exposed_comm_fraction"] = (ms_dev - ms_local) / ms_dev
    tr.close()

    # ---- the same-precision arm: the same step with tf32 tensor-core math on fp32 activations (the precision of
    # ---- the reference's fp32 model under torch's TF32 convolutions), same mode, same batch, same lease
    if args.dtype == "bf16" and not args.no_tf32:
        import argparse as _ap
        a32 = _ap.Namespace(**{**vars(args), "dtype": "tf32"})
        t = Trainer(native_config(a32, mode, world, rank, base_port + 10, strong_main))
        for i in range(3):
            t.backend.run_staged(t.backend.stage_inputs(*host_batches[i % 8]))
        steps = min(args.steps, 50)
        ms = time_replays(t.backend, steps)
        out["same_precision"] = {"dtype": "tf32", "ms_per_step": ms, "images_per_s": cfg.global_batch / (ms * 1e-3),
                                 "steps": steps, "roofline": roofline(specs, B, ms, 4)}
        t.close()

    # ---- the reference's own operating point: global batch 256 split over the ranks, every sync mode
    if not args.no_strong:
        strong = {}
        gb = args.global_batch
        if world == 1:
            if not strong_main and B == gb:
                strong["part1"] = {"ms_per_step": ms_dev, "images_per_s": gb / (ms_dev * 1e-3)}
        elif gb % world == 0:
            steps = min(args.steps, 50)
            for k, m in enumerate(("part3", "part2b", "part2a")):
                if strong_main and m == mode:
                    strong[m] = {"ms_per_step": ms_dev, "images_per_s": gb / (ms_dev * 1e-3)}
                    continue
                t = Trainer(native_config(args, m, world, rank, base_port + 1 + k, True))
                hb = make_batches(gb // world, rank, 4)
                for i in range(3):
                    t.backend.run_staged(t.backend.stage_inputs(*hb[i % 4]))
                ms = time_replays(t.backend, steps)
                strong[m] = {"ms_per_step": ms, "images_per_s": gb / (ms * 1e-3), "steps": steps}
                t.close()
        out["strong"] = {"global_batch": gb, "per_gpu_batch": gb // world, **strong}
    return out if rank == 0 else {}


def run_reference(args) -> dict:
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "baseline"))
    import ref_harness
    why = ref_harness.available()
    if why is not None:
        return {"impl": "reference", "unavailable": why}
    import torch.distributed as dist
    world = env_int("WORLD_SIZE", 1)
    rank = env_int("RANK", 0)
    local = env_int("LOCAL_RANK", 0)
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}; launch with torch.distributed.run for N>1")
    # one GPU: the tutorial's single-process script (part1), as for the native arm
    mode = args.mode if world > 1 else "part1"
    B = args.per_gpu_batch if args.scaling == "weak" else args.global_batch // world
    torch.cuda.set_device(local)
    sampler = ClockSampler(local)

    def max_ranks(v: float) -> float:
        if world == 1:
            return v
        t = torch.tensor([v], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    try:
        dev = ref_harness.run_reference(mode, rank, world, local, B, args.steps, max(args.warmup, 3), host_inputs=False,
                                        on_timed_region=sampler.start, on_timed_region_end=sampler.stop)
        e2e = None
        if not args.no_e2e:
            e2e = ref_harness.run_reference(mode, rank, world, local, B, args.steps, 3, host_inputs=True)
    except Exception as e:  # the reference arm must never take the benchmark down
        return {"impl": "reference", "unavailable": f"{type(e).__name__}: {e}"[:300]}

    ms_dev = max_ranks(dev["ms_per_step"])
    gb = B * world
    out = {
        "metric": "VGG-11 CIFAR-10 images/sec", "impl": "reference", "value": gb / (ms_dev * 1e-3), "unit": "images/s",
        "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_dev,
        "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "fp32 (torch default, TF32 conv)",
        "data": "synthetic (random CIFAR-10-shaped tensors, random-init weights)",
        "config": {"model": "VGG-11", "global_batch": gb, "per_gpu_batch": B, "seq_len": None, "image": "3x32x32",
                   "mode": mode, "parallelism": f"dp{world}", "optimizer": "SGD lr0.1 m0.9 wd1e-4", "backend": "nccl",
                   "script": "/".join(ref_harness.SCRIPTS[mode]),
                   "note": "unmodified train_model() and model.py from baseline/_ref; device global set to cuda"},
        "gpu_launches": 0, "clocks": sampler.summary(),
    }
    if e2e is not None:
        ms = max_ranks(e2e["ms_per_step"])
        out["e2e"] = {"value": gb / (ms * 1e-3), "unit": "images/s", "ms_per_step": ms,
                      "h2d_bytes_per_step": e2e["h2d_bytes_per_step"], "d2h_bytes_per_step": 4.0 / 20}
    # ---- with the tutorial's own input pipeline in the loop (CIFAR10 dataset class, PIL transforms,
    # ---- DataLoader(num_workers=2, pin_memory=True)) instead of ready-made tensors
    if not args.no_loader:
        try:
            steps = min(args.steps, 50)
            r = ref_harness.run_reference(mode, rank, world, local, B, steps, 5, host_inputs=True, real_loader=True)
            ms = max_ranks(r["ms_per_step"])
            out["e2e_with_loader"] = {"value": gb / (ms * 1e-3), "unit": "images/s", "ms_per_step": ms, "steps": steps,
                                      "loader": "torchvision CIFAR10 + transforms, DataLoader(num_workers=2)"}
        except Exception as e:
            out["e2e_with_loader"] = {"unavailable": f"{type(e).__name__}: {e}"[:200]}
    # ---- the reference's own operating point: global batch 256 split over the ranks, every sync mode
    if not args.no_strong:
        strong = {}
        g256 = args.global_batch
        if world == 1:
            if args.scaling == "weak" and B == g256:
                strong["part1"] = {"ms_per_step": ms_dev, "images_per_s": g256 / (ms_dev * 1e-3)}
        elif g256 % world == 0:
            steps = min(args.steps, 50)
            for m in ("part3", "part2b", "part2a"):
                try:
                    r = ref_harness.run_reference(m, rank, world, local, g256 // world, steps, 3, host_inputs=False)
                    ms = max_ranks(r["ms_per_step"])
                    strong[m] = {"ms_per_step": ms, "images_per_s": g256 / (ms * 1e-3), "steps": steps}
                except Exception as e:
                    strong[m] = {"unavailable": f"{type(e).__name__}: {e}"[:200]}
        out["strong"] = {"global_batch": g256, "per_gpu_batch": g256 // world, **strong}
    if dist.is_initialized():
        dist.destroy_process_group()
    return out if rank == 0 else {}


def main():
    args = parse_args()
    if not torch.cuda.is_available():
        print(json.dumps({"impl": args.impl, "unavailable": "no CUDA device visible"}))
        return 0
    out = run_reference(args) if args.impl == "reference" else run_native(args)
    if out:
        print(json.dumps(out), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
