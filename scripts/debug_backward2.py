import copy, sys, torch, torch.nn.functional as F
sys.path.insert(0, "tests")
from emulated_reference import emulated_forward
from cs744_pytorch_distributed_tutorial_b200.models.vgg import VGG
torch.backends.cudnn.allow_tf32 = False; torch.backends.cuda.matmul.allow_tf32 = False
DEV = "cuda"
def rel(a, b): return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-12)).item()
def cos(a, b): return F.cosine_similarity(a.float().flatten(), b.float().flatten(), dim=0).item()
for B in (32, 256):
    torch.manual_seed(5000)
    ref = VGG("VGG11").to(DEV); emu = copy.deepcopy(ref); model = copy.deepcopy(ref); model.enable_native(max_batch=B)
    x = torch.randn(B, 3, 32, 32, device=DEV); t = torch.randint(0, 10, (B,), device=DEV)
    ref.train(); emu.train(); model.train()
    F.cross_entropy(ref(x), t).backward()
    le = emulated_forward(emu, x); F.cross_entropy(le, t).backward()
    out = model(x); F.cross_entropy(out, t).backward()
    torch.cuda.synchronize()
    print("B", B, "logits vs emu", rel(out, le))
    rp, ep = dict(ref.named_parameters()), dict(emu.named_parameters())
    for n, p in model.named_parameters():
        if ep[n].grad.norm() < 1e-5: continue
        print(f"  {n:18s} vs emu rel {rel(p.grad, ep[n].grad):.4f} cos {cos(p.grad, ep[n].grad):.5f} | vs fp32 rel {rel(p.grad, rp[n].grad):.4f} cos {cos(p.grad, rp[n].grad):.5f} | emu vs fp32 rel {rel(ep[n].grad, rp[n].grad):.4f}")
# does the stock model also blow up on a repeated batch at lr 0.1?
torch.manual_seed(5000)
ref = VGG("VGG11").to(DEV); opt = torch.optim.SGD(ref.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
x = torch.randn(32, 3, 32, 32, device=DEV); t = torch.randint(0, 10, (32,), device=DEV)
for i in range(6):
    opt.zero_grad(); l = F.cross_entropy(ref(x), t); l.backward(); opt.step(); print("stock torch step", i, l.item())
