"""Smallest full training step (batch 32, eager, no autotune) for compute-sanitizer runs."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cs744_pytorch_distributed_tutorial_b200.models.vgg import VGG11
from cs744_pytorch_distributed_tutorial_b200.ops import native as K
torch.manual_seed(0)
m = VGG11().cuda()
nv = m.enable_native(max_batch=32, autotune=False)
x = torch.randn(32, 3, 32, 32, device="cuda"); t = torch.randint(0, 10, (32,), device="cuda")
for _ in range(2):
    nv.forward(x, True); nv.loss_fwd_bwd(t); nv.backward()
    K.sgd_momentum(nv.flat_params, nv.flat_grads, nv.flat_momentum, 0.01, 0.9, 1e-4); nv.refresh_shadows()
# also the split-K and CTA-pair variants
b = nv.blocks[-1]
K.conv_igemm(b.x_in[:32], b.wf, b.bias, b.y[:32], stats=b.stats, block_n=128, split_k=4, workspace=nv.splitk_ws)
b4 = nv.blocks[4]
K.conv_igemm(b4.x_in[:32], b4.wf, b4.bias, b4.y[:32], stats=b4.stats, block_n=128, cta_pair=True)
torch.cuda.synchronize()
print("loss", float(nv.loss))
