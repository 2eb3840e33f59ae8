"""Acceptance tests of the fused paths the executor takes by default: the data gradient whose epilogue also
reduces the BatchNorm backward of the block below and the one-launch classifier head.  The optional
weight-gradient -> reduce-scatter kernel (B200DP_WGRAD_PUSH=1) is checked on >= 2 GPUs.
"""
import os

import pytest
import torch

pytestmark = [pytest.mark.gpu]

DEV = "cuda"


def _native():
    from cs744_pytorch_distributed_tutorial_b200.ops import native
    return native


# (B, H, W, C_dy = Cout of the conv, C_prev = Cin of the conv = channels of the block below)
DGRAD_SHAPES = [
    (32, 16, 16, 128, 64),
    (32, 8, 8, 256, 128),
    (32, 8, 8, 256, 256),
    (32, 4, 4, 512, 256),
    (32, 4, 4, 512, 512),
    (32, 2, 2, 512, 512),
    (30, 2, 2, 512, 512),    # last 128-row tile only partly valid
]


def _dgrad_cfg_ok(shape, cfg):
    B, H, W, C, Cp = shape
    block_n, split_k, pair = cfg
    return (not block_n or Cp % block_n == 0) and not (pair and ((B * H * W + 127) // 128) % 2)


# (N tile, split-K, CTA pair): only what the shape admits (N tile divides the channels, even tile count for pairs)
@pytest.mark.parametrize("shape,cfg", [
    (s, c) for s in DGRAD_SHAPES
    for c in [(0, 1, False), (32, 1, False), (64, 1, False), (128, 1, False), (256, 1, False), (64, 1, True),
              (128, 1, True), (256, 1, True), (64, 3, False), (128, 4, True)] if _dgrad_cfg_ok(s, c)])
def test_dgrad_with_bn_reduce_matches_the_two_pass_path(shape, cfg):
    """Same dL/da bits as the plain dgrad kernel, and the same sums as bn_bwd_reduce_from_out on it."""
    native = _native()
    from cs744_pytorch_distributed_tutorial_b200.ops._lib import call, ptr, stream_ptr
    B, H, W, C, Cp = shape
    block_n, split_k, pair = cfg
    g = torch.Generator(device=DEV).manual_seed(11)
    dy = torch.randn(B, H, W, C, device=DEV, generator=g).bfloat16()
    wd = (torch.randn(Cp, 9 * C, device=DEV, generator=g) / (3 * C ** 0.5)).bfloat16()
    a = torch.relu(torch.randn(B, H, W, Cp, device=DEV, generator=g)).bfloat16()   # about half zeros
    gamma = torch.rand(Cp, device=DEV, generator=g) + 0.5
    gamma[0] = 0.0                                                                   # guarded division
    beta = torch.randn(Cp, device=DEV, generator=g) * 0.1
    ws = torch.zeros(B * H * W * Cp, device=DEV)

    ref = torch.empty(B, H, W, Cp, device=DEV, dtype=torch.bfloat16)
    native.conv_igemm(dy, wd, None, ref, block_n=block_n, split_k=split_k, cta_pair=pair, workspace=ws)
    sums_ref = torch.zeros(2 * Cp, device=DEV)
    call("b200dp_bn_relu_pool_bwd_reduce_from_out", ptr(a), ptr(ref), ptr(gamma), ptr(beta), ptr(sums_ref),
         B * H * W, Cp, stream_ptr())

    out = torch.full_like(ref, float("nan"))
    sums = torch.zeros(2 * Cp, device=DEV)
    native.conv_dgrad_bn_reduce(dy, wd, out, a, gamma, beta, sums, block_n=block_n, split_k=split_k,
                                workspace=ws, cta_pair=pair)
    torch.cuda.synchronize()
    if split_k == 1:
        assert torch.equal(out, ref)
    else:   # fp32 atomics: order-dependent last bit before the bf16 rounding
        assert ((out.float() - ref.float()).norm() / ref.float().norm()).item() < 2e-3
    assert float(ws.abs().max()) == 0.0          # split-K workspace handed back clean
    scale = sums_ref.abs().max().item() + 1e-6
    assert (sums - sums_ref).abs().max().item() < 2e-3 * scale, (sums[:8], sums_ref[:8])
    assert sums[Cp].item() == 0.0                # gamma == 0 channel contributes nothing to S2


@pytest.mark.parametrize("B", [256, 64, 37, 1])
def test_fused_head_matches_the_three_kernel_path(B):
    """b200dp_head_fwd_bwd vs linear_fwd + cross_entropy + linear_bwd, and vs torch fp32."""
    native = _native()
    Fd, O = 512, 10
    g = torch.Generator(device=DEV).manual_seed(3)
    x = torch.randn(B, 1, 1, Fd, device=DEV, generator=g).bfloat16()
    w = torch.randn(O, Fd, device=DEV, generator=g) * 0.05
    b = torch.randn(O, device=DEV, generator=g) * 0.1
    t = torch.randint(0, O, (B,), device=DEV, generator=g)
    # reference: the three validated kernels
    logits0 = torch.empty(B, O, device=DEV); dl0 = torch.empty(B, O, device=DEV)
    loss0 = torch.zeros((), device=DEV); corr0 = torch.zeros((), device=DEV, dtype=torch.int32)
    dw0 = torch.zeros(O, Fd, device=DEV); db0 = torch.zeros(O, device=DEV)
    dx0 = torch.empty_like(x)
    native.linear_fwd(x, w, b, logits0)
    native.cross_entropy(logits0, t, loss0, dl0, corr0, 1.0)
    native.linear_bwd(x, w, dl0, dw0, db0, dx0)
    for rep in range(2):      # twice: the block counter must have been reset by the first launch
        logits = torch.full((B, O), float("nan"), device=DEV); dl = torch.full((B, O), float("nan"), device=DEV)
        loss = torch.full((), float("nan"), device=DEV); corr = torch.full((), -7, device=DEV, dtype=torch.int32)
        dw = torch.zeros(O, Fd, device=DEV); db = torch.zeros(O, device=DEV)
        dx = torch.full_like(x, float("nan"))
        native.head_fwd_bwd(x, w, b, t, logits, loss, corr, dl, dx, dw, db, 1.0)
        torch.cuda.synchronize()
        torch.testing.assert_close(logits, logits0, rtol=1e-5, atol=1e-5)
        torch.testing.assert_close(dl, dl0, rtol=1e-4, atol=1e-7)
        torch.testing.assert_close(loss, loss0, rtol=1e-5, atol=1e-6)
        assert int(corr) == int(corr0)
        torch.testing.assert_close(dx.float(), dx0.float(), rtol=2e-2, atol=1e-6)
        torch.testing.assert_close(dw, dw0, rtol=1e-4, atol=1e-6)
        torch.testing.assert_close(db, db0, rtol=1e-4, atol=1e-7)
    # and the plain fp32 definition
    ref_logits = x.float().view(B, Fd) @ w.t() + b
    torch.testing.assert_close(logits, ref_logits, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(loss, torch.nn.functional.cross_entropy(ref_logits, t), rtol=1e-4, atol=1e-5)


def test_training_step_with_fused_head_matches_the_default():
    from cs744_pytorch_distributed_tutorial_b200.models.vgg import VGG
    torch.manual_seed(0)
    x = torch.randn(64, 3, 32, 32, device=DEV)
    y = torch.randint(0, 10, (64,), device=DEV)
    res = []
    torch.manual_seed(1)
    model = VGG("VGG11").to(DEV)
    # ONE executor for both flavours: a second one would autotune its own tile shapes, and a different
    # accumulation order of the bf16 products alone moves the loss by ~2e-3 (seen on hardware)
    nv = model.enable_native(max_batch=64)
    for defer in (False, True):
        nv.defer_head = defer
        for _ in range(2):      # second pass: flags and gradient clearing must have been reset
            nv.forward(x, training=True)
            nv.loss_fwd_bwd(y)
            nv.backward()
        torch.cuda.synchronize()
        res.append((float(nv.loss), int(nv.correct), nv.flat_grads.clone()))
    (l0, c0, g0), (l1, c1, g1) = res
    # Two runs of the SAME flavour already differ by ~1e-3 in the loss on hardware: the BatchNorm statistics are
    # accumulated with float atomics, and a last-bit change of a mean moves bf16 roundings downstream.  The exact
    # comparison of the head itself is test_fused_head_matches_the_three_kernel_path above.
    assert abs(l0 - l1) < 5e-3 and abs(c0 - c1) <= 1
    assert torch.nn.functional.cosine_similarity(g0, g1, dim=0).item() > 0.97   # measured 0.986 between equal runs


# ---------------------------------------------------------------------------------------------------
# >= 2 GPUs: weight-gradient GEMM -> reduce-scatter in one kernel (B200DP_WGRAD_PUSH=1)
# ---------------------------------------------------------------------------------------------------
def _push_worker(rank, world, push, port, outdir):
    import os as _os
    _os.environ["B200DP_WGRAD_PUSH"] = "1" if push else "0"
    _os.environ["LOCAL_RANK"] = str(rank)
    from cs744_pytorch_distributed_tutorial_b200.config import TrainConfig
    from cs744_pytorch_distributed_tutorial_b200.engine import Trainer
    torch.cuda.set_device(rank)
    cfg = TrainConfig(mode="part3", master_ip="127.0.0.1", num_nodes=world, rank=rank, port=port, device="cuda",
                      per_rank_batch=32, eval=False, lr=0.01, timeout_s=60, bucket_plan="layer")   # one conv per bucket
    tr = Trainer(cfg)
    nv = tr.backend.native
    assert (nv.wgrad_push is not None) == bool(push)
    init = nv.flat_params.cpu()
    g = torch.Generator().manual_seed(300 + rank)
    losses, p1 = [], None
    for _ in range(3):
        x = torch.randn(32, 3, 32, 32, generator=g).pin_memory()
        y = torch.randint(0, 10, (32,), generator=g).pin_memory()
        losses.append(float(tr.train_step(x, y).item()))
        if p1 is None:
            torch.cuda.synchronize()
            p1 = nv.flat_params.cpu()
    torch.cuda.synchronize()
    torch.save({"params": nv.flat_params.cpu(), "p1": p1, "init": init, "losses": losses},
               _os.path.join(outdir, f"push{int(push)}_{rank}.pt"))
    tr.close()


@pytest.mark.multigpu
def test_wgrad_push_reduce_scatter_matches_the_bucket_allreduce():
    import sys
    import tempfile
    import torch.multiprocessing as mp
    sys.path.insert(0, os.path.dirname(__file__))
    from helpers import free_port
    n = torch.cuda.device_count()
    world = 8 if n >= 8 else 4 if n >= 4 else 2
    res = {}
    with tempfile.TemporaryDirectory() as d:
        for push in (False, True):
            mp.spawn(_push_worker, args=(world, push, free_port(), d), nprocs=world, join=True)
            res[push] = [torch.load(os.path.join(d, f"push{int(push)}_{r}.pt")) for r in range(world)]
    for push, ranks in res.items():
        for r in ranks[1:]:
            assert torch.equal(ranks[0]["params"], r["params"]), f"push={push}: replicas differ"
    a, b = res[False][0], res[True][0]
    assert torch.equal(a["init"], b["init"])
    assert abs(a["losses"][0] - b["losses"][0]) < 1e-2
    # the FIRST update is compared: it is the mean gradient of identical inputs on identical weights, so the two
    # paths differ only by bf16 accumulation order (tile choices) and float atomics; later steps of two
    # independent runs drift apart on their own (ReLU / max-pool decisions flip), whichever path is taken
    da, db = a["p1"] - a["init"], b["p1"] - b["init"]
    cos = torch.nn.functional.cosine_similarity(da, db, dim=0).item()
    assert cos > 0.97, cos          # two runs of the SAME path measure 0.986 (bf16 + float atomics)
    assert abs(db.norm().item() / da.norm().item() - 1) < 0.03


def test_classifier_forward_runs_on_the_tcgen05_gemm_kernel_too():
    """The 512 -> 10 linear layer as a taps=1 implicit GEMM on the tensor cores (weights padded to 32
    rows): same kernel as layer 0.  Not the default: the whole layer is 2.6 MFLOP, less than the fixed
    cost of a TMEM / TMA pipeline, and the GEMM kernel stores bf16 logits."""
    native = _native()
    B, Fd, O = 256, 512, 10
    g = torch.Generator(device=DEV).manual_seed(4)
    x = torch.randn(B, 1, 1, Fd, device=DEV, generator=g).bfloat16()
    w = torch.zeros(32, Fd, device=DEV)
    w[:O] = torch.randn(O, Fd, device=DEV, generator=g) * 0.05
    b = torch.zeros(32, device=DEV)
    b[:O] = torch.randn(O, device=DEV, generator=g) * 0.1
    out = torch.empty(B, 1, 1, 32, device=DEV, dtype=torch.bfloat16)
    native.conv_igemm(x, w.bfloat16().contiguous(), b, out, taps=1, block_n=32)
    torch.cuda.synchronize()
    ref = x.float().view(B, Fd) @ w.bfloat16().float().t() + b
    torch.testing.assert_close(out.float().view(B, 32)[:, :O], ref[:, :O], rtol=2e-2, atol=2e-2)
    assert float(out.float().view(B, 32)[:, O:].abs().max()) == 0.0
