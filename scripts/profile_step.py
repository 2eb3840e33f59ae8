"""Run a few eager (no CUDA graph) training steps for ncu / per-kernel timing.

    python scripts/profile_step.py [--mode part3] [--batch 256] [--steps 2] [--events]

--events prints a per-kernel-family breakdown measured with CUDA events around each native call
(serialised, so it over-counts launch gaps; use for shares, not absolutes)."""
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
a = ap.parse_args()

cfg = TrainConfig(mode=a.mode, per_rank_batch=a.batch, device="cuda", eval=False, cuda_graph=False,
                  single_rank_collectives=True)
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
else:
    # with `ncu --profile-from-start off` only these steps are captured (not the autotuner's runs)
    torch.cuda.profiler.start()
    for _ in range(a.steps):
        tr.train_step(x, y)
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
tr.close()
