"""Numerics of every single-GPU sm_100a kernel against a plain PyTorch fp32 reference
of the same op.  Shapes are the VGG-11 layer shapes of SURVEY.md 2.7.
"""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

DEV = "cuda"


def _native():
    from cs744_pytorch_distributed_tutorial_b200.ops import native
    return native


def _rel_err(a: torch.Tensor, b: torch.Tensor) -> float:
    return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-12)).item()


@pytest.fixture(autouse=True)
def _fp32_reference():
    old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old


# (B, H, W, Cin, Cout)
CONV_SHAPES = [
    (32, 16, 16, 64, 128),   # layers.4
    (32, 8, 8, 128, 256),    # layers.8
    (32, 8, 8, 256, 256),    # layers.11
    (32, 4, 4, 256, 512),    # layers.15
    (32, 4, 4, 512, 512),    # layers.18
    (32, 2, 2, 512, 512),    # layers.22 / 25
    (64, 2, 2, 512, 512),
    (2, 32, 32, 64, 64),     # W = 32 geometry
]


def _conv_ref(x_nhwc, w_krsc, bias):
    x = x_nhwc.float().permute(0, 3, 1, 2)
    w = w_krsc.float().permute(0, 3, 1, 2)
    y = F.conv2d(x, w, bias, padding=1)
    return y.permute(0, 2, 3, 1).contiguous()


# only the N tiles that divide Cout (0 = let the launcher choose)
@pytest.mark.parametrize("shape,block_n", [(s, n) for s in CONV_SHAPES for n in (0, 32, 64, 128, 256)
                                           if not n or s[4] % n == 0])
def test_conv_fwd(shape, block_n):
    native = _native()
    B, H, W, Cin, Cout = shape
    g = torch.Generator(device=DEV).manual_seed(1)
    x = torch.randn(B, H, W, Cin, device=DEV, generator=g).bfloat16()
    w = (torch.randn(Cout, 3, 3, Cin, device=DEV, generator=g) / (3 * Cin ** 0.5)).bfloat16()
    bias = torch.randn(Cout, device=DEV, generator=g)
    out = torch.empty(B, H, W, Cout, device=DEV, dtype=torch.bfloat16)
    stats = torch.zeros(2, Cout, device=DEV)
    native.conv_igemm(x, w.view(Cout, 9 * Cin), bias, out, stats=stats, block_n=block_n)
    torch.cuda.synchronize()
    ref = _conv_ref(x, w, bias)
    err = _rel_err(out, ref)
    assert err < 6e-3, f"conv fwd rel err {err}"
    # fused BatchNorm statistics are those of the stored bf16 values
    o32 = out.float().view(-1, Cout)
    torch.testing.assert_close(stats[0], o32.sum(0), rtol=2e-3, atol=2e-2)
    torch.testing.assert_close(stats[1], (o32 * o32).sum(0), rtol=2e-3, atol=2e-2)


@pytest.mark.parametrize("shape", CONV_SHAPES)
def test_conv_dgrad(shape):
    native = _native()
    B, H, W, Cin, Cout = shape
    g = torch.Generator(device=DEV).manual_seed(2)
    dy = torch.randn(B, H, W, Cout, device=DEV, generator=g).bfloat16()
    w = (torch.randn(Cout, 3, 3, Cin, device=DEV, generator=g) / (3 * Cout ** 0.5)).bfloat16()
    master = w.float().contiguous()
    wf = torch.empty(Cout, 9 * Cin, device=DEV, dtype=torch.bfloat16)
    wd = torch.empty(Cin, 9 * Cout, device=DEV, dtype=torch.bfloat16)
    native.prep_conv_weights(master, wf, wd, Cout, Cin, 9, 9 * Cin)
    assert torch.equal(wf.view(Cout, 3, 3, Cin), w)
    wd_ref = w.flip(1, 2).permute(3, 1, 2, 0).contiguous()  # [Cin][r'][s'][Cout]
    assert torch.equal(wd.view(Cin, 3, 3, Cout), wd_ref)
    dx = torch.empty(B, H, W, Cin, device=DEV, dtype=torch.bfloat16)
    native.conv_igemm(dy, wd, None, dx)
    torch.cuda.synchronize()
    ref = F.conv_transpose2d(dy.float().permute(0, 3, 1, 2), w.float().permute(0, 3, 1, 2), padding=1)
    err = _rel_err(dx, ref.permute(0, 2, 3, 1))
    assert err < 6e-3, f"conv dgrad rel err {err}"


@pytest.mark.parametrize("shape", CONV_SHAPES)
@pytest.mark.parametrize("split_k", [0, 1, 3])
def test_conv_wgrad(shape, split_k):
    native = _native()
    B, H, W, Cin, Cout = shape
    g = torch.Generator(device=DEV).manual_seed(3)
    x = torch.randn(B, H, W, Cin, device=DEV, generator=g).bfloat16()
    dy = (torch.randn(B, H, W, Cout, device=DEV, generator=g) / (B * H * W) ** 0.5).bfloat16()
    dw = torch.zeros(Cout, 3, 3, Cin, device=DEV)
    native.conv_wgrad(dy, x, dw, split_k=split_k)
    torch.cuda.synchronize()
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(False)
    wref = torch.zeros(Cout, Cin, 3, 3, device=DEV, requires_grad=True)
    F.conv2d(xr, wref, None, padding=1).backward(dy.float().permute(0, 3, 1, 2))
    ref = wref.grad.permute(0, 2, 3, 1)
    err = _rel_err(dw, ref)
    assert err < 2e-3, f"conv wgrad rel err {err}"


def test_im2col_that_leaves_the_zero_half_alone():
    """``upper_half_zeroed``: only columns 0..31 of the bf16 patch rows are written (the executor's matrix is
    allocated as zeros and columns 32..63 never change)."""
    native = _native()
    B, H, W = 5, 32, 32
    g = torch.Generator(device=DEV).manual_seed(14)
    x = torch.randn(B, 3, H, W, device=DEV, generator=g)
    full = torch.empty(B, H, W, 64, device=DEV, dtype=torch.bfloat16)
    native.im2col3x3_c3(x, full)
    lo = torch.full((B, H, W, 64), 7.0, device=DEV, dtype=torch.bfloat16)
    native.im2col3x3_c3(x, lo, upper_half_zeroed=True)
    torch.cuda.synchronize()
    assert torch.equal(lo[..., :32], full[..., :32])
    assert bool((lo[..., 32:] == 7.0).all())          # untouched


def test_layer0_as_gemm():
    """3-channel first conv = patch gather + K=64 GEMM; its wgrad writes only the 27 real columns."""
    native = _native()
    B, H, W, Cout = 8, 32, 32, 64
    g = torch.Generator(device=DEV).manual_seed(4)
    x = torch.randn(B, 3, H, W, device=DEV, generator=g)
    w = torch.randn(Cout, 3, 3, 3, device=DEV, generator=g) / 5
    bias = torch.randn(Cout, device=DEV, generator=g)
    a0 = torch.empty(B, H, W, 64, device=DEV, dtype=torch.bfloat16)
    native.im2col3x3_c3(x, a0)
    patches = F.unfold(x, 3, padding=1).view(B, 3, 9, H * W).permute(0, 3, 2, 1).reshape(B, H, W, 27)
    assert torch.equal(a0[..., :27], patches.bfloat16())
    assert a0[..., 27:].abs().max().item() == 0
    wf = torch.empty(Cout, 64, device=DEV, dtype=torch.bfloat16)
    native.prep_conv_weights(w.contiguous(), wf, None, Cout, 3, 9, 64)
    assert torch.equal(wf[:, :27], w.view(Cout, 27).bfloat16()) and wf[:, 27:].abs().max().item() == 0
    out = torch.empty(B, H, W, Cout, device=DEV, dtype=torch.bfloat16)
    native.conv_igemm(a0, wf, bias, out, taps=1)
    torch.cuda.synchronize()
    ref = F.conv2d(x.bfloat16().float(), w.bfloat16().float().permute(0, 3, 1, 2), bias, padding=1)
    assert _rel_err(out, ref.permute(0, 2, 3, 1)) < 6e-3
    dy = (torch.randn(B, H, W, Cout, device=DEV, generator=g) / 50).bfloat16()
    dw = torch.zeros(Cout, 27, device=DEV)
    native.conv_wgrad(dy, a0, dw, taps=1, cin_valid=27)
    torch.cuda.synchronize()
    ref_dw = dy.float().view(-1, Cout).t() @ a0.float().view(-1, 64)[:, :27]
    assert _rel_err(dw, ref_dw) < 2e-3


# the VGG shapes run as bulk-copy pipelines (csrc/bn_stream.cu); 24 channels (3 chunks do not divide a 256-thread
# stage) and 6x6 pooled (3 x 16 items per pooled row) take the register-file kernels
@pytest.mark.parametrize("C,H,W,pool", [(64, 32, 32, True), (128, 16, 16, True), (256, 8, 8, False),
                                        (512, 4, 4, True), (512, 2, 2, False), (512, 2, 2, True),
                                        (24, 8, 8, False), (24, 8, 8, True), (128, 6, 6, True)])
@pytest.mark.parametrize("B", [16, 37, 256])
def test_bn_relu_pool_fwd_bwd(C, H, W, pool, B):
    native = _native()
    g = torch.Generator(device=DEV).manual_seed(5)
    y = (torch.randn(B, H, W, C, device=DEV, generator=g) * 1.5 + 0.3).bfloat16()
    gamma = torch.rand(C, device=DEV, generator=g) + 0.5
    beta = torch.randn(C, device=DEV, generator=g) * 0.2
    rm = torch.randn(C, device=DEV, generator=g) * 0.1
    rv = torch.rand(C, device=DEV, generator=g) + 0.5
    nbt = torch.zeros((), device=DEV, dtype=torch.int64)
    rm0, rv0 = rm.clone(), rv.clone()
    stats = torch.zeros(2, C, device=DEV)
    native.bn_stats(y, stats)
    Ho, Wo = (H // 2, W // 2) if pool else (H, W)
    out = torch.empty(B, Ho, Wo, C, device=DEV, dtype=torch.bfloat16)
    mean = torch.empty(C, device=DEV)
    invstd = torch.empty(C, device=DEV)
    native.bn_relu_pool_fwd(y, stats, gamma, beta, rm, rv, nbt, mean, invstd, out, pool, True)
    torch.cuda.synchronize()

    yr = y.float().permute(0, 3, 1, 2).requires_grad_(True)
    gr = gamma.clone().requires_grad_(True)
    br = beta.clone().requires_grad_(True)
    rm_ref, rv_ref = rm0.clone(), rv0.clone()
    z = F.batch_norm(yr, rm_ref, rv_ref, gr, br, training=True, momentum=0.1, eps=1e-5)
    a = F.relu(z)
    o = F.max_pool2d(a, 2, 2) if pool else a
    torch.testing.assert_close(out.float(), o.permute(0, 2, 3, 1).detach(), rtol=1e-2, atol=1e-2)
    torch.testing.assert_close(rm, rm_ref, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(rv, rv_ref, rtol=1e-4, atol=1e-5)
    assert nbt.item() == 1

    dout = torch.randn(B, Ho, Wo, C, device=DEV, generator=g).bfloat16()
    o.backward(dout.float().permute(0, 3, 1, 2))
    sums = torch.zeros(2, C, device=DEV)
    dy = torch.empty_like(y)
    dgamma = torch.empty(C, device=DEV)
    dbeta = torch.empty(C, device=DEV)
    native.bn_relu_pool_bwd(y, dout, mean, invstd, gamma, beta, sums, dy, dgamma, dbeta, pool)
    torch.cuda.synchronize()
    assert _rel_err(dgamma, gr.grad) < 2e-3
    assert _rel_err(dbeta, br.grad) < 2e-3
    assert _rel_err(dy, yr.grad.permute(0, 2, 3, 1)) < 8e-3

    # the same backward with the reduction taken from the stored block output (the executor's default)
    sums2 = torch.zeros(2, C, device=DEV)
    dy2 = torch.empty_like(y)
    dgamma2 = torch.empty(C, device=DEV)
    dbeta2 = torch.empty(C, device=DEV)
    native.bn_relu_pool_bwd(y, dout, mean, invstd, gamma, beta, sums2, dy2, dgamma2, dbeta2, pool, out=out)
    torch.cuda.synchronize()
    assert _rel_err(dbeta2, br.grad) < 2e-3
    assert _rel_err(dgamma2, gr.grad) < 5e-3       # xhat recovered from the bf16-rounded output
    assert _rel_err(dy2, yr.grad.permute(0, 2, 3, 1)) < 8e-3

    # eval mode uses the running statistics
    out_e = torch.empty_like(out)
    native.bn_relu_pool_fwd(y, None, gamma, beta, rm, rv, None, None, None, out_e, pool, False)
    ze = F.relu(F.batch_norm(y.float().permute(0, 3, 1, 2), rm, rv, gamma, beta, training=False, eps=1e-5))
    oe = F.max_pool2d(ze, 2, 2) if pool else ze
    torch.testing.assert_close(out_e.float(), oe.permute(0, 2, 3, 1), rtol=1e-2, atol=1e-2)


@pytest.mark.parametrize("B", [32, 256, 300])
def test_linear_cross_entropy(B):
    native = _native()
    g = torch.Generator(device=DEV).manual_seed(6)
    x = torch.randn(B, 512, device=DEV, generator=g).bfloat16()
    w = torch.randn(10, 512, device=DEV, generator=g) / 20
    b = torch.randn(10, device=DEV, generator=g)
    t = torch.randint(0, 10, (B,), device=DEV, generator=g)
    logits = torch.empty(B, 10, device=DEV)
    loss = torch.empty((), device=DEV)
    dlogits = torch.empty(B, 10, device=DEV)
    correct = torch.zeros((), device=DEV, dtype=torch.int32)
    native.linear_fwd(x, w, b, logits)
    native.cross_entropy(logits, t, loss, dlogits, correct)
    dw = torch.empty_like(w)
    db = torch.empty_like(b)
    dx = torch.empty_like(x)
    native.linear_bwd(x, w, dlogits, dw, db, dx)
    torch.cuda.synchronize()
    xr = x.float().requires_grad_(True)
    wr = w.clone().requires_grad_(True)
    br = b.clone().requires_grad_(True)
    lr = F.linear(xr, wr, br)
    ls = F.cross_entropy(lr, t)
    ls.backward()
    torch.testing.assert_close(logits, lr.detach(), rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(loss, ls.detach(), rtol=1e-5, atol=1e-5)
    assert correct.item() == (lr.argmax(1) == t).sum().item()
    torch.testing.assert_close(dw, wr.grad, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(db, br.grad, rtol=1e-4, atol=1e-5)
    assert _rel_err(dx, xr.grad) < 5e-3


@pytest.mark.parametrize("n", [10, 4099, 9231114])
def test_sgd_momentum_matches_torch(n):
    native = _native()
    g = torch.Generator(device=DEV).manual_seed(7)
    p = torch.randn(n, device=DEV, generator=g)
    buf = torch.zeros(n, device=DEV)
    pr = p.clone().requires_grad_(True)
    opt = torch.optim.SGD([pr], lr=0.1, momentum=0.9, weight_decay=1e-4)
    for _ in range(3):
        grad = torch.randn(n, device=DEV, generator=g)
        native.sgd_momentum(p, grad, buf, 0.1, 0.9, 1e-4)
        pr.grad = grad.clone()
        opt.step()
    torch.cuda.synchronize()
    torch.testing.assert_close(p, pr.detach(), rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("shape,block_n,split_k", [((64, 2, 2, 512, 512), 128, 4), ((64, 2, 2, 512, 512), 64, 9),
                                                   ((32, 4, 4, 256, 512), 256, 3), ((32, 8, 8, 128, 256), 128, 2)])
def test_conv_fwd_split_k(shape, block_n, split_k):
    """K loop split across CTAs: partial tiles meet in an fp32 workspace, the finalize kernel adds the
    bias, rounds, takes the statistics and leaves the workspace zero for the next call."""
    native = _native()
    B, H, W, Cin, Cout = shape
    g = torch.Generator(device=DEV).manual_seed(9)
    x = torch.randn(B, H, W, Cin, device=DEV, generator=g).bfloat16()
    w = (torch.randn(Cout, 3, 3, Cin, device=DEV, generator=g) / (3 * Cin ** 0.5)).bfloat16()
    bias = torch.randn(Cout, device=DEV, generator=g)
    ws = torch.zeros(B * H * W * Cout, device=DEV)
    ref = _conv_ref(x, w, bias)
    for _ in range(2):   # twice: the second call relies on the workspace having been re-zeroed
        out = torch.empty(B, H, W, Cout, device=DEV, dtype=torch.bfloat16)
        stats = torch.zeros(2, Cout, device=DEV)
        native.conv_igemm(x, w.view(Cout, 9 * Cin), bias, out, stats=stats, block_n=block_n, split_k=split_k,
                          workspace=ws)
        torch.cuda.synchronize()
        assert _rel_err(out, ref) < 6e-3
        o32 = out.float().view(-1, Cout)
        torch.testing.assert_close(stats[0], o32.sum(0), rtol=2e-3, atol=2e-2)
        torch.testing.assert_close(stats[1], (o32 * o32).sum(0), rtol=2e-3, atol=2e-2)
        assert ws.abs().max().item() == 0


# the CTA-pair kernel needs an even number of 128-pixel tiles and an N tile that divides Cout
@pytest.mark.parametrize("shape,block_n,split_k", [
    (s, n, k) for s in [(32, 16, 16, 64, 128), (32, 8, 8, 256, 256), (32, 4, 4, 512, 512), (64, 2, 2, 512, 512),
                        (2, 32, 32, 64, 64)]
    for n, k in [(64, 1), (128, 1), (256, 1), (128, 3)]
    if s[4] % n == 0 and ((s[0] * s[1] * s[2]) // 128) % 2 == 0])
def test_conv_fwd_cta_pair(shape, block_n, split_k):
    """tcgen05.mma.cta_group::2: two CTAs share one 256-row tile, each staging half of the weights."""
    native = _native()
    B, H, W, Cin, Cout = shape
    g = torch.Generator(device=DEV).manual_seed(10)
    x = torch.randn(B, H, W, Cin, device=DEV, generator=g).bfloat16()
    w = (torch.randn(Cout, 3, 3, Cin, device=DEV, generator=g) / (3 * Cin ** 0.5)).bfloat16()
    bias = torch.randn(Cout, device=DEV, generator=g)
    out = torch.empty(B, H, W, Cout, device=DEV, dtype=torch.bfloat16)
    stats = torch.zeros(2, Cout, device=DEV)
    ws = torch.zeros(out.numel(), device=DEV) if split_k > 1 else None
    native.conv_igemm(x, w.view(Cout, 9 * Cin), bias, out, stats=stats, block_n=block_n, split_k=split_k,
                      workspace=ws, cta_pair=True)
    torch.cuda.synchronize()
    ref = _conv_ref(x, w, bias)
    assert _rel_err(out, ref) < 6e-3
    o32 = out.float().view(-1, Cout)
    torch.testing.assert_close(stats[0], o32.sum(0), rtol=2e-3, atol=2e-2)
    torch.testing.assert_close(stats[1], (o32 * o32).sum(0), rtol=2e-3, atol=2e-2)
