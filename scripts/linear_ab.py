"""A/B of the classifier forward (256 x 512 -> 10): CUDA-core kernels vs the tcgen05 GEMM kernel.

    python scripts/linear_ab.py > gpurun_out/linear_ab.txt

Each candidate is captured as 20 back-to-back launches in a CUDA graph and replayed; the time per launch is the
CUDA-event time of a replay / 20 (best of 20 replays), so launch latency is amortised the way it is in the step
graph.  The tcgen05 path is the taps=1 implicit GEMM of layer 0 with the weight padded to 32 rows; to be a full
replacement it would still need a loss kernel behind it (its logits are rounded to the activation type).
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cs744_pytorch_distributed_tutorial_b200.ops import native as K  # noqa: E402

DEV = torch.device("cuda", 0)
torch.cuda.set_device(DEV)
B, F, O = 256, 512, 10
g = torch.Generator(device=DEV).manual_seed(0)
x = torch.randn(B, F, device=DEV, generator=g).bfloat16()
w = torch.randn(O, F, device=DEV, generator=g) * 0.05
b = torch.randn(O, device=DEV, generator=g) * 0.1
t = torch.randint(0, O, (B,), device=DEV, generator=g)
logits = torch.empty(B, O, device=DEV)
loss = torch.zeros(1, device=DEV)
correct = torch.zeros(1, device=DEV, dtype=torch.int32)
dlogits = torch.empty(B, O, device=DEV)
dx = torch.empty(B, F, device=DEV, dtype=torch.bfloat16)
dw = torch.zeros(O, F, device=DEV)
db = torch.zeros(O, device=DEV)
w32 = torch.zeros(32, F, device=DEV)
w32[:O] = w
w32 = w32.bfloat16().contiguous()
b32 = torch.zeros(32, device=DEV)
b32[:O] = b
out32 = torch.empty(B, 1, 1, 32, device=DEV, dtype=torch.bfloat16)
x4 = x.view(B, 1, 1, F)

CANDIDATES = {
    "linear_fwd (CUDA cores, warp per sample)": lambda: K.linear_fwd(x, w, b, logits),
    "conv_igemm taps=1, N=32 (tcgen05, 2 CTAs)": lambda: K.conv_igemm(x4, w32, b32, out32, taps=1, block_n=32),
    "linear_fwd + cross_entropy + linear_bwd (3 launches)": lambda: (
        K.linear_fwd(x, w, b, logits), K.cross_entropy(logits, t, loss, dlogits, correct, 1.0),
        K.linear_bwd(x, w, dlogits, dw, db, dx)),
    "head_fwd_bwd (1 launch, the default)": lambda: K.head_fwd_bwd(x, w, b, t, logits, loss, correct, dlogits, dx, dw,
                                                                     db, 1.0),
}


def time_graph(fn, reps=20, replays=20):
    s = torch.cuda.Stream(DEV)
    with torch.cuda.stream(s):
        for _ in range(3):
            fn()
        s.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=s):
            for _ in range(reps):
                fn()
        best = float("inf")
        for _ in range(replays):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(s)
            graph.replay()
            e1.record(s)
            e1.synchronize()
            best = min(best, e0.elapsed_time(e1) * 1e3 / reps)
    return best


print(f"classifier {B} x {F} -> {O}, bf16 features; us per call, 20 calls per graph replay, best of 20 replays")
for name, fn in CANDIDATES.items():
    print(f"{time_graph(fn):8.2f} us  {name}")
ref = x.float() @ w.t() + b
K.linear_fwd(x, w, b, logits)
K.conv_igemm(x4, w32, b32, out32, taps=1, block_n=32)
torch.cuda.synchronize()
print(f"max |logit error| vs fp32 matmul: CUDA cores {float((logits - ref).abs().max()):.2e}, "
      f"tcgen05 (bf16 weights, bf16 logits) {float((out32.view(B, 32)[:, :O].float() - ref).abs().max()):.2e}")
