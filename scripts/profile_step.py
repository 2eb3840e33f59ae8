"""Run a few eager (no CUDA graph) training steps for ncu / per-kernel timing.

    python scripts/profile_step.py [--mode part3] [--batch 256] [--steps 2] [--events | --select]

--events prints a per-kernel-family breakdown measured with CUDA events around each native call
(serialised, so it over-counts launch gaps; use for shares, not absolutes).

--select (for `ncu --profile-from-start off`): the profiler range is opened only around a fixed choice of native
calls of ONE step -- one of each kernel family at its best and worst shape -- which keeps a `--set full` report
with imported source under the 64 MiB that travels back from the GPU box."""
import argparse
import os
import sys
import collections

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cs744_pytorch_distributed_tutorial_b200.config import TrainConfig  # noqa: E402
from cs744_pytorch_distributed_tutorial_b200.engine import Trainer  # noqa: E402
from cs744_pytorch_distributed_tutorial_b200.ops import _lib  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--mode", default="part3")
ap.add_argument("--batch", type=int, default=256)
ap.add_argument("--steps", type=int, default=2)
ap.add_argument("--warmup", type=int, default=3)
ap.add_argument("--events", action="store_true")
ap.add_argument("--select", action="store_true")
ap.add_argument("--dtype", default="bf16")
a = ap.parse_args()

cfg = TrainConfig(mode=a.mode, per_rank_batch=a.batch, device="cuda", eval=False, cuda_graph=False,
                  single_rank_collectives=True, dtype=a.dtype)
tr = Trainer(cfg)
be = tr.backend
x = torch.randn(a.batch, 3, 32, 32).pin_memory()
y = torch.randint(0, 10, (a.batch,)).pin_memory()
for _ in range(a.warmup):
    tr.train_step(x, y)
torch.cuda.synchronize()

if a.events:
    records = []
    orig = _lib.call

    def timed_call(name, *args):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        orig(name, *args)
        e1.record()
        records.append((name, args, e0, e1))
    import cs744_pytorch_distributed_tutorial_b200.ops.native as N
    import cs744_pytorch_distributed_tutorial_b200.parallel.comm as C
    N.call = timed_call
    C.call = timed_call
    for _ in range(a.steps):
        be._step_body()
    torch.cuda.synchronize()
    N.call = orig
    C.call = orig
    agg = collections.OrderedDict()
    for name, args, e0, e1 in records:
        key = name
        if name in ("b200dp_conv_igemm", "b200dp_conv_wgrad"):
            ints = [v for v in args if isinstance(v, int) and not isinstance(v, bool) and 0 < v < 100000][:6]
            key = f"{name} {ints}"
        agg.setdefault(key, []).append(e0.elapsed_time(e1) * 1e3)
    total = 0.0
    for k, v in agg.items():
        per_step = sum(v) / a.steps
        total += per_step
        print(f"{per_step:9.1f} us/step  x{len(v) // a.steps:<3d} {k}")
    print(f"{total:9.1f} us/step  total (serialised, event-bracketed)")
elif a.select:
    # occurrence index of each call inside one step; forward calls run blocks 0..7, backward calls blocks 7..0
    SELECT = {
        "b200dp_im2col3x3_c3_ex2": {0},
        "b200dp_conv_igemm_ex": {0, 1, 3, 7},                 # layer 0 GEMM, 64->128@16x16, 256->256@8x8, 512->512@2x2
        "b200dp_bn_relu_pool_fwd_ex": {0, 2},
        "b200dp_head_fwd_bwd_ex": {0},
        "b200dp_bn_relu_pool_bwd_apply_ex": {5, 7},            # blocks 2, 0
        "b200dp_conv_dgrad_bn_reduce_ex": {4, 6},              # blocks 3, 1
        "b200dp_conv_wgrad_ex": {4, 6, 7},                     # blocks 3, 1, 0
        "b200dp_reducer_bucket_ready": {0}, "b200dp_sgd_momentum": {0},
        "b200dp_prep_conv_weights_batched_ex": {0},
    }
    seen = collections.Counter()
    orig = _lib.call

    def selected_call(name, *args):
        k = seen[name]
        seen[name] += 1
        if k in SELECT.get(name, ()):
            torch.cuda.synchronize()
            torch.cuda.profiler.start()
            try:
                return orig(name, *args)
            finally:
                torch.cuda.synchronize()
                torch.cuda.profiler.stop()
        return orig(name, *args)
    import cs744_pytorch_distributed_tutorial_b200.ops.native as N
    import cs744_pytorch_distributed_tutorial_b200.parallel.comm as C
    N.call = selected_call
    C.call = selected_call
    tr.train_step(x, y)
    torch.cuda.synchronize()
    N.call = orig
    C.call = orig
    print("calls of the step:", dict(seen))
else:
    # with `ncu --profile-from-start off` only these steps are captured (not the autotuner's runs)
    torch.cuda.profiler.start()
    for _ in range(a.steps):
        tr.train_step(x, y)
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
tr.close()
