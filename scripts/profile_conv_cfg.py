"""Graph-timed (kernel-only) comparison of conv tile configurations at the VGG-11 shapes."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cs744_pytorch_distributed_tutorial_b200.ops import native as K
from cs744_pytorch_distributed_tutorial_b200.models.native_vgg import _time_captured, _igemm_candidates
B = 256
shapes = [(16, 64, 128), (8, 128, 256), (8, 256, 256), (4, 256, 512), (4, 512, 512), (2, 512, 512)]
ws = torch.zeros(B * 16 * 16 * 128, device="cuda")
s = torch.cuda.Stream()
with torch.cuda.stream(s):
    for hw, cin, cout in shapes:
        x = torch.randn(B, hw, hw, cin, device="cuda").bfloat16()
        w = torch.randn(cout, 9 * cin, device="cuda").bfloat16()
        bias = torch.randn(cout, device="cuda"); y = torch.empty(B, hw, hw, cout, device="cuda", dtype=torch.bfloat16)
        stats = torch.zeros(2, cout, device="cuda")
        flops = 2.0 * B * hw * hw * cout * 9 * cin
        res = []
        for c in _igemm_candidates((B * hw * hw + 127) // 128, cout, 9 * cin // 64):
            t = _time_captured(lambda c=c: K.conv_igemm(x, w, bias, y, stats=stats, block_n=c[0], split_k=c[1], cta_pair=c[2], workspace=ws))
            res.append((t, c))
        res.sort()
        print(f"fwd [{B},{hw},{hw},{cin}->{cout}]  " + "  ".join(f"{c}:{t:.1f}us" for t, c in res[:6]) + f"   best {flops / res[0][0] / 1e6:.0f} TFLOP/s")
