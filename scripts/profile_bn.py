import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cs744_pytorch_distributed_tutorial_b200.ops import native as K
B, hw, C, pool = 256, 32, 64, True
y = torch.randn(B, hw, hw, C, device="cuda").bfloat16(); out = torch.empty(B, hw // 2, hw // 2, C, device="cuda", dtype=torch.bfloat16)
dout = torch.randn_like(out); dy = torch.empty_like(y)
stats = torch.zeros(2, C, device="cuda"); sums = torch.zeros(2, C, device="cuda"); K.bn_stats(y, stats)
g = torch.ones(C, device="cuda"); b = torch.zeros(C, device="cuda"); rm = torch.zeros(C, device="cuda"); rv = torch.ones(C, device="cuda")
nbt = torch.zeros((), device="cuda", dtype=torch.int64); mean = torch.empty(C, device="cuda"); invstd = torch.empty(C, device="cuda")
dg = torch.empty(C, device="cuda"); db = torch.empty(C, device="cuda")
for _ in range(3):
    K.bn_relu_pool_fwd(y, stats, g, b, rm, rv, nbt, mean, invstd, out, pool, True)
    sums.zero_()
    K.bn_relu_pool_bwd(y, dout, mean, invstd, g, b, sums, dy, dg, db, pool, fused=False)
torch.cuda.synchronize()
