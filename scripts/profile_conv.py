"""Time (CUDA events, L2 flushed between iterations) or profile single conv kernels.

    python scripts/profile_conv.py --kind fwd|dgrad|wgrad --shape B,H,W,Cin,Cout [--block-n N] [--iters 20]
"""
import argparse, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cs744_pytorch_distributed_tutorial_b200.ops import native as K

ap = argparse.ArgumentParser()
ap.add_argument("--kind", default="fwd")
ap.add_argument("--shape", default="256,8,8,256,256")
ap.add_argument("--block-n", type=int, default=0)
ap.add_argument("--split-k", type=int, default=0)
ap.add_argument("--iters", type=int, default=20)
ap.add_argument("--all", action="store_true", help="sweep the VGG-11 layer shapes at B=256")
a = ap.parse_args()

VGG11 = [(256, 16, 16, 64, 128), (256, 8, 8, 128, 256), (256, 8, 8, 256, 256), (256, 4, 4, 256, 512),
         (256, 4, 4, 512, 512), (256, 2, 2, 512, 512)]
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def bench(kind, shape, block_n, split_k, iters):
    B, H, W, Cin, Cout = shape
    x = torch.randn(B, H, W, Cin, device="cuda").bfloat16()
    dy = torch.randn(B, H, W, Cout, device="cuda").bfloat16()
    w = torch.randn(Cout, 9 * Cin, device="cuda").bfloat16()
    wd = torch.randn(Cin, 9 * Cout, device="cuda").bfloat16()
    bias = torch.randn(Cout, device="cuda")
    y = torch.empty(B, H, W, Cout, device="cuda", dtype=torch.bfloat16)
    dx = torch.empty(B, H, W, Cin, device="cuda", dtype=torch.bfloat16)
    dw = torch.zeros(Cout, 9, Cin, device="cuda")
    stats = torch.zeros(2, Cout, device="cuda")
    if kind == "fwd":
        fn = lambda: K.conv_igemm(x, w, bias, y, stats=stats, block_n=block_n)
    elif kind == "dgrad":
        fn = lambda: K.conv_igemm(dy, wd, None, dx, block_n=block_n)
    else:
        fn = lambda: K.conv_wgrad(dy, x, dw, split_k=split_k)
    for _ in range(3):
        fn()
    times = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1) * 1e3)
    times.sort()
    med = times[len(times) // 2]
    flops = 2.0 * B * H * W * Cout * 9 * Cin
    print(f"{kind:5s} {str(shape):28s} block_n={block_n:<3d} split_k={split_k:<2d} median {med:7.1f} us  "
          f"{flops / med / 1e6:7.1f} TFLOP/s  min {times[0]:.1f}")
    return med


if a.all:
    for kind in ("fwd", "dgrad", "wgrad"):
        for s in VGG11:
            if kind == "fwd":
                for bn in (32, 64, 128, 256):
                    if s[4] % bn == 0:
                        bench(kind, s, bn, 0, a.iters)
            elif kind == "dgrad":
                for bn in (32, 64, 128, 256):
                    if s[3] % bn == 0:
                        bench(kind, s, bn, 0, a.iters)
            else:
                for sk in (0, 1, 2, 4, 8):
                    bench(kind, s, 0, sk, a.iters)
else:
    bench(a.kind, tuple(int(v) for v in a.shape.split(",")), a.block_n, a.split_k, a.iters)
