"""CUDA-event timing (L2 flushed between iterations) of the bandwidth-bound kernels at VGG-11 shapes."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cs744_pytorch_distributed_tutorial_b200.ops import native as K

B = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 256
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
HBM = 6548.5e9


WARM = "--warm" in sys.argv   # back-to-back launches, no L2 flush: kernel-only time, L2-resident inputs


def timeit(name, fn, nbytes, iters=10):
    for _ in range(3):
        fn()
    if WARM:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(50):
            fn()
        e1.record(); torch.cuda.synchronize()
        med = e0.elapsed_time(e1) * 1e3 / 50
        print(f"{name:44s} {med:7.1f} us  {nbytes / 1e6:7.1f} MB  {nbytes / med / 1e6:6.2f} TB/s (back-to-back, L2 warm)")
        return med
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    med = ts[len(ts) // 2]
    print(f"{name:44s} {med:7.1f} us  {nbytes / 1e6:7.1f} MB  {nbytes / med / 1e6:6.2f} TB/s ({100 * nbytes / med / 1e-6 / HBM:4.0f}% of measured HBM)")
    return med


tot = 0.0
layers = [(32, 64, True), (16, 128, True), (8, 256, False), (8, 256, True), (4, 512, False), (4, 512, True),
          (2, 512, False), (2, 512, True)]
for hw, C, pool in layers:
    y = torch.randn(B, hw, hw, C, device="cuda").bfloat16()
    ho = hw // 2 if pool else hw
    out = torch.empty(B, ho, ho, C, device="cuda", dtype=torch.bfloat16)
    dout = torch.randn(B, ho, ho, C, device="cuda").bfloat16()
    dy = torch.empty_like(y)
    stats = torch.zeros(2, C, device="cuda"); sums = torch.zeros(2, C, device="cuda")
    K.bn_stats(y, stats)
    g = torch.ones(C, device="cuda"); b = torch.zeros(C, device="cuda")
    rm = torch.zeros(C, device="cuda"); rv = torch.ones(C, device="cuda"); nbt = torch.zeros((), device="cuda", dtype=torch.int64)
    mean = torch.empty(C, device="cuda"); invstd = torch.empty(C, device="cuda")
    dg = torch.empty(C, device="cuda"); db = torch.empty(C, device="cuda")
    tag = f"[{B},{hw},{hw},{C}] pool={int(pool)}"
    tot += timeit("bn_relu_pool_fwd " + tag, lambda: K.bn_relu_pool_fwd(y, stats, g, b, rm, rv, nbt, mean, invstd, out, pool, True),
                  y.numel() * 2 + out.numel() * 2)
    from cs744_pytorch_distributed_tutorial_b200.ops._lib import call, ptr, stream_ptr
    red = lambda: call("b200dp_bn_relu_pool_bwd_reduce", ptr(y), ptr(dout), ptr(mean), ptr(invstd), ptr(g), ptr(b), ptr(sums), B, hw, hw, C, int(pool), stream_ptr())
    app = lambda: call("b200dp_bn_relu_pool_bwd_apply", ptr(y), ptr(dout), ptr(mean), ptr(invstd), ptr(g), ptr(b), ptr(sums), ptr(dy), ptr(dg), ptr(db), B, hw, hw, C, int(pool), stream_ptr())
    tot += timeit("bn_bwd_reduce    " + tag, red, y.numel() * 2 + dout.numel() * 2)
    tot += timeit("bn_bwd_apply     " + tag, app, y.numel() * 4 + dout.numel() * 2)
x = torch.randn(B, 3, 32, 32, device="cuda"); a0 = torch.empty(B, 32, 32, 64, device="cuda", dtype=torch.bfloat16)
tot += timeit("im2col3x3_c3", lambda: K.im2col3x3_c3(x, a0), x.numel() * 4 + a0.numel() * 2)
feat = torch.randn(B, 512, device="cuda").bfloat16(); w = torch.randn(10, 512, device="cuda"); bb = torch.zeros(10, device="cuda")
dl = torch.randn(B, 10, device="cuda"); dw = torch.empty_like(w); dbb = torch.empty(10, device="cuda"); dx = torch.empty_like(feat)
tot += timeit("linear_bwd", lambda: K.linear_bwd(feat, w, dl, dw, dbb, dx), feat.numel() * 4)
n = 9_232_384
p = torch.randn(n, device="cuda"); gr = torch.randn(n, device="cuda"); m = torch.zeros(n, device="cuda")
tot += timeit("sgd_momentum (9.2M)", lambda: K.sgd_momentum(p, gr, m, 0.1, 0.9, 1e-4), n * 4 * 5)
cfg = [(128, 64), (256, 128), (256, 256), (512, 256), (512, 512), (512, 512), (512, 512)]
ls = []
for co, ci in cfg:
    ls.append((torch.randn(co, 9, ci, device="cuda"), torch.empty(co, 9 * ci, device="cuda", dtype=torch.bfloat16),
               torch.empty(ci, 9 * co, device="cuda", dtype=torch.bfloat16), co, ci, 9, 9 * ci))
prep = K.BatchedWeightPrep(ls)
nb = sum(l[0].numel() * 8 for l in ls)
tot += timeit("prep_conv_weights_batched (7 layers)", prep, nb)
print(f"sum of medians: {tot:.1f} us (each includes ~launch latency of an eager call)")
