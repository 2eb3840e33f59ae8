"""In-graph timeline of one training step, from CUPTI through torch.profiler (nsys is not installed).

    python scripts/timeline_step.py [--mode part1|part3] [--batch 256] [--steps 3] [--out gpurun_out/timeline]

Unlike the `ncu` launch list (kernels serialised, L2 flushed between them) this records the kernels of the
REPLAYED CUDA graph with their real start times and overlap: which kernels run next to each other on the
compute, weight-gradient and reducer streams, where the gaps are, and how long the critical path is.
Prints, for the last recorded step: one line per kernel (start offset, duration, stream), the busy time
per stream, the wall time of the step and the idle gaps on the compute stream larger than 1 us.  Also
writes a chrome trace (`<out>.json`) and the table (`<out>.txt`).

Numbers are taken under a profiler: use them to understand the step, never as benchmark results.
With torchrun the script traces rank 0 only (all ranks run the steps).
"""
import argparse
import os
import sys

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cs744_pytorch_distributed_tutorial_b200.config import TrainConfig  # noqa: E402
from cs744_pytorch_distributed_tutorial_b200.engine import Trainer  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--mode", default="part1")
ap.add_argument("--batch", type=int, default=256)
ap.add_argument("--steps", type=int, default=3)
ap.add_argument("--warmup", type=int, default=5)
ap.add_argument("--out", default="gpurun_out/timeline")
ap.add_argument("--single-rank-collectives", action="store_true")
a = ap.parse_args()

rank = int(os.environ.get("RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
cfg = TrainConfig(mode=a.mode, per_rank_batch=a.batch, device="cuda", eval=False, num_nodes=world, rank=rank,
                  master_ip=os.environ.get("MASTER_ADDR", "127.0.0.1"),
                  port=int(os.environ.get("MASTER_PORT", "29500")) + 17,
                  single_rank_collectives=a.single_rank_collectives)
tr = Trainer(cfg)
x = torch.randn(a.batch, 3, 32, 32).pin_memory()
y = torch.randint(0, 10, (a.batch,)).pin_memory()
for _ in range(a.warmup):
    tr.train_step(x, y)
torch.cuda.synchronize()

with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(a.steps):
        tr.train_step(x, y)
        torch.cuda.synchronize()      # keeps the steps apart in the trace
torch.cuda.synchronize()

if rank == 0:
    os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
    prof.export_chrome_trace(a.out + ".json")
    # kernel records: name, start (us), duration (us), stream
    ks = []
    for ev in prof.events():
        if getattr(ev, "device_type", None) is not None and "CUDA" in str(ev.device_type):
            rng = ev.time_range      # device events: start / end of the kernel or copy on the GPU, in us
            ks.append((rng.start, rng.end - rng.start, ev.name, getattr(ev, "device_resource_id", -1)))
    ks.sort()
    if not ks:
        print("no device activity recorded (is CUPTI available?)")
    else:
        # split into steps at gaps > 200 us (the synchronize between steps), keep the last one
        steps, cur = [], [ks[0]]
        for prev, k in zip(ks, ks[1:]):
            if k[0] - (prev[0] + prev[1]) > 200:
                steps.append(cur)
                cur = []
            cur.append(k)
        steps.append(cur)
        # the H2D copies of the next batch form their own tiny groups: take the largest group
        step = max(steps, key=len)
        t0 = step[0][0]
        end = max(s + d for s, d, _, _ in step)
        lines = [f"# {a.mode}, batch {a.batch}, {len(step)} device activities, wall {end - t0:.1f} us"]
        busy = {}
        for s, d, name, stream in step:
            short = name.replace("void ", "").replace("b200dp::", "").split("(")[0][:60]
            lines.append(f"{s - t0:9.1f} us  +{d:7.1f} us  stream {stream:<4} {short}")
            busy[stream] = busy.get(stream, 0.0) + d
        lines.append("# busy time per stream: " + ", ".join(f"{k}: {v:.1f} us" for k, v in sorted(busy.items())))
        main = max(busy, key=busy.get)
        on_main = sorted((s, d, n) for s, d, n, st in step if st == main)
        gaps = [(b[0] - (p[0] + p[1]), p[2], b[2]) for p, b in zip(on_main, on_main[1:]) if b[0] - (p[0] + p[1]) > 1.0]
        lines.append(f"# gaps > 1 us on stream {main}: total {sum(g for g, _, _ in gaps):.1f} us")
        for g, before, after in sorted(gaps, reverse=True)[:25]:
            lines.append(f"#   {g:6.1f} us between {before.split('(')[0][-40:]} and {after.split('(')[0][-40:]}")
        text = "\n".join(lines)
        print(text)
        with open(a.out + ".txt", "w") as f:
            f.write(text + "\n")
tr.close()
