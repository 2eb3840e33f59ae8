"""Compare per-block intermediates of the native executor with the fp32 torch model."""
import copy, torch, torch.nn.functional as F
from cs744_pytorch_distributed_tutorial_b200.models.vgg import VGG
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
DEV = "cuda"
def rel(a, b): return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-12)).item()
torch.manual_seed(5000)
B = 32
ref = VGG("VGG11").to(DEV); model = copy.deepcopy(ref); model.enable_native(max_batch=B)
nv = model.native
x = torch.randn(B, 3, 32, 32, device=DEV); t = torch.randint(0, 10, (B,), device=DEV)
acts, grads = {}, {}
def fwd_hook(name):
    def h(m, i, o):
        acts[name] = o.detach().clone()
        o.register_hook(lambda g: grads.__setitem__(name, g.detach().clone()))
    return h
for i, m in enumerate(ref.layers):
    if isinstance(m, (torch.nn.Conv2d, torch.nn.MaxPool2d)) or (isinstance(m, torch.nn.ReLU)):
        m.register_forward_hook(fwd_hook(i))
# ReLU is inplace: hook BN instead for pre-relu
ref.train(); model.train()
out_ref = ref(x); F.cross_entropy(out_ref, t).backward()
nv.forward(x, True); nv.loss_fwd_bwd(t); nv.backward()
torch.cuda.synchronize()
for bi, b in enumerate(nv.blocks):
    s = b.spec
    y_ref = acts[s.conv_idx].permute(0, 2, 3, 1)
    dy_ref = grads[s.conv_idx].permute(0, 2, 3, 1)
    print(f"block {bi} conv {s.conv_idx}: y {rel(b.y, y_ref):.4f}  dy {rel(b.dy, dy_ref):.4f}  |dy_ref| {dy_ref.norm().item():.3e} |dy| {b.dy.float().norm().item():.3e}", end="")
    out_idx = s.conv_idx + (3 if s.pool else 2)
    o_ref = acts[out_idx].permute(0, 2, 3, 1); do_ref = grads[out_idx].permute(0, 2, 3, 1)
    print(f"  out {rel(b.out, o_ref):.4f} dout {rel(b.dout, do_ref):.4f}")
rp = dict(ref.named_parameters())
for n, p in model.named_parameters():
    print(n, f"{rel(p.grad, rp[n].grad):.4f}", f"{rp[n].grad.norm().item():.3e}")
