import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cs744_pytorch_distributed_tutorial_b200.ops import native as K
from cs744_pytorch_distributed_tutorial_b200.models.native_vgg import _time_captured
B = 256
x = torch.randn(B, 3, 32, 32, device="cuda"); w = torch.randn(64, 27, device="cuda") / 5; bias = torch.randn(64, device="cuda")
y = torch.empty(B, 32, 32, 64, device="cuda", dtype=torch.bfloat16); stats = torch.zeros(2, 64, device="cuda")
dy = torch.randn(B, 32, 32, 64, device="cuda").bfloat16(); dw = torch.zeros(64, 27, device="cuda")
a0 = torch.empty(B, 32, 32, 64, device="cuda", dtype=torch.bfloat16); wf = torch.zeros(64, 64, device="cuda", dtype=torch.bfloat16)
s = torch.cuda.Stream()
with torch.cuda.stream(s):
    print("direct fwd        %.1f us" % _time_captured(lambda: K.conv3x3_c3_fwd(x, w, bias, y, stats)))
    print("direct fwd nostat %.1f us" % _time_captured(lambda: K.conv3x3_c3_fwd(x, w, bias, y, None)))
    print("direct wgrad      %.1f us" % _time_captured(lambda: K.conv3x3_c3_wgrad(dy, x, dw)))
    print("im2col            %.1f us" % _time_captured(lambda: K.im2col3x3_c3(x, a0)))
    print("gemm fwd (K=64)   %.1f us" % _time_captured(lambda: K.conv_igemm(a0, wf, bias, y, stats=stats, taps=1)))
    print("gemm wgrad        %.1f us" % _time_captured(lambda: K.conv_wgrad(dy, a0, dw, taps=1, cin_valid=27)))
