"""The same-precision arm: ``tcgen05.mma kind::tf32`` kernels on fp32 activations, against plain PyTorch fp32
ops of the same layer, FROM FP32 INPUTS (nothing is pre-rounded: the bound covers the whole precision gap to
the reference's fp32 model, ref: master/part1/model.py:18-25, which torch runs with TF32 convolutions on a GPU).
A tf32 operand keeps 10 mantissa bits (relative rounding error 2^-11 per element, averaged down by the K-long
fp32 accumulation), so 1e-3 relative L2 is the acceptance bound for every GEMM.
"""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

DEV = "cuda"
TOL = 1e-3


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


# (B, H, W, Cin, Cout): the VGG-11 layer shapes, a W = 32 geometry, and a ragged batch (partly valid last tile)
CONV_SHAPES = [
    (32, 16, 16, 64, 128), (32, 8, 8, 128, 256), (32, 8, 8, 256, 256), (32, 4, 4, 256, 512), (32, 4, 4, 512, 512),
    (32, 2, 2, 512, 512), (256, 8, 8, 128, 256), (2, 32, 32, 64, 64), (2, 32, 32, 32, 64), (30, 2, 2, 512, 512),
]


def _conv_ref(x_nhwc, w_krsc, bias):
    y = F.conv2d(x_nhwc.permute(0, 3, 1, 2), w_krsc.permute(0, 3, 1, 2), bias, padding=1)
    return y.permute(0, 2, 3, 1).contiguous()


@pytest.mark.parametrize("shape", CONV_SHAPES)
@pytest.mark.parametrize("cfg", [(0, 1, False), (32, 1, False), (64, 1, False), (128, 1, False), (256, 1, False),
                                 (128, 1, True), (256, 1, True), (64, 3, False), (128, 2, True)])
def test_conv_fwd_tf32(shape, cfg):
    native = _native()
    B, H, W, Cin, Cout = shape
    block_n, split_k, pair = cfg
    m_tiles = (B * H * W + 127) // 128
    if (block_n and Cout % block_n) or (pair and m_tiles % 2) or split_k > 9 * Cin // 32:
        pytest.skip("configuration does not apply to this shape")
    g = torch.Generator(device=DEV).manual_seed(1)
    x = torch.randn(B, H, W, Cin, device=DEV, generator=g)
    w = torch.randn(Cout, 3, 3, Cin, device=DEV, generator=g) / (3 * Cin ** 0.5)
    bias = torch.randn(Cout, device=DEV, generator=g)
    out = torch.full((B, H, W, Cout), float("nan"), device=DEV)
    stats = torch.zeros(2, Cout, device=DEV)
    ws = torch.zeros(B * H * W * Cout, device=DEV)
    native.conv_igemm(x, w.view(Cout, 9 * Cin), bias, out, stats=stats, block_n=block_n, split_k=split_k, cta_pair=pair,
                      workspace=ws)
    torch.cuda.synchronize()
    ref = _conv_ref(x, w, bias)
    err = _rel_err(out, ref)
    assert err < TOL, f"tf32 conv fwd rel err {err}"
    o = out.view(-1, Cout)
    torch.testing.assert_close(stats[0], o.sum(0), rtol=1e-3, atol=2e-2)
    torch.testing.assert_close(stats[1], (o * o).sum(0), rtol=1e-3, atol=2e-2)
    assert float(ws.abs().max()) == 0.0          # the split-K workspace is handed back clean


@pytest.mark.parametrize("shape", CONV_SHAPES)
@pytest.mark.parametrize("cfg", [(0, 1, False), (64, 1, False), (128, 1, True), (64, 3, False)])
def test_conv_dgrad_with_bn_reduce_tf32(shape, cfg):
    """Data gradient (tap-flipped transposed fp32 weights) and, in the same launch, the BatchNorm-backward sums of
    the block below, against their definitions in fp32."""
    native = _native()
    B, H, W, Cin, Cout = shape
    block_n, split_k, pair = cfg
    m_tiles = (B * H * W + 127) // 128
    if (block_n and Cin % block_n) or (pair and m_tiles % 2) or Cout % 32:
        pytest.skip("configuration does not apply to this shape")
    g = torch.Generator(device=DEV).manual_seed(2)
    dy = torch.randn(B, H, W, Cout, device=DEV, generator=g)
    w = torch.randn(Cout, 3, 3, Cin, device=DEV, generator=g) / (3 * Cout ** 0.5)
    wd = torch.empty(Cin, 9 * Cout, device=DEV)
    native.prep_conv_weights(w.contiguous(), None, wd, Cout, Cin, 9, 9 * Cin)
    assert torch.equal(wd.view(Cin, 3, 3, Cout), w.flip(1, 2).permute(3, 1, 2, 0).contiguous())
    a_prev = torch.relu(torch.randn(B, H, W, Cin, device=DEV, generator=g))
    gamma = torch.rand(Cin, device=DEV, generator=g) + 0.5
    beta = torch.randn(Cin, device=DEV, generator=g) * 0.1
    sums = torch.zeros(2, Cin, device=DEV)
    dx = torch.full((B, H, W, Cin), float("nan"), device=DEV)
    ws = torch.zeros(B * H * W * Cin, device=DEV)
    native.conv_dgrad_bn_reduce(dy, wd, dx, a_prev, gamma, beta, sums, block_n=block_n, split_k=split_k, cta_pair=pair,
                                workspace=ws)
    torch.cuda.synchronize()
    ref = F.conv_transpose2d(dy.permute(0, 3, 1, 2), w.permute(0, 3, 1, 2), padding=1).permute(0, 2, 3, 1)
    err = _rel_err(dx, ref)
    assert err < TOL, f"tf32 conv dgrad rel err {err}"
    d = torch.where(a_prev > 0, dx, torch.zeros_like(dx)).view(-1, Cin)
    s1 = d.sum(0)
    s2 = (d * ((a_prev.view(-1, Cin) - beta) / gamma)).sum(0)
    scale = max(float(s1.abs().max()), float(s2.abs().max()))
    assert float((sums[0] - s1).abs().max()) < 1e-3 * scale
    assert float((sums[1] - s2).abs().max()) < 1e-3 * scale


@pytest.mark.parametrize("shape", CONV_SHAPES)
@pytest.mark.parametrize("split_k", [0, 1, 3])
def test_conv_wgrad_tf32(shape, split_k):
    native = _native()
    B, H, W, Cin, Cout = shape
    if Cout % 128 and Cout != 64:
        pytest.skip("M tile")
    g = torch.Generator(device=DEV).manual_seed(3)
    x = torch.randn(B, H, W, Cin, device=DEV, generator=g)
    dy = torch.randn(B, H, W, Cout, device=DEV, generator=g) / (B * H * W) ** 0.5
    dw = torch.zeros(Cout, 3, 3, Cin, device=DEV)
    native.conv_wgrad(dy, x, dw, split_k=split_k)
    torch.cuda.synchronize()
    wref = torch.zeros(Cout, Cin, 3, 3, device=DEV, requires_grad=True)
    F.conv2d(x.permute(0, 3, 1, 2), wref, None, padding=1).backward(dy.permute(0, 3, 1, 2))
    err = _rel_err(dw, wref.grad.permute(0, 2, 3, 1))
    assert err < TOL, f"tf32 conv wgrad rel err {err}"
