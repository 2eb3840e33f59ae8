"""Kernels written at the end of round 1 after the GPU budget was spent: compiled, SASS-checked, not yet
run on hardware.  They are off by default in the executor; these tests are their acceptance checks and
run only with ``B200DP_EXPERIMENTAL=1`` so that an unvalidated path cannot turn the suite red.

    B200DP_EXPERIMENTAL=1 python -m pytest tests/test_gpu_experimental.py -m gpu -x -q
"""
import os

import pytest
import torch

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(os.environ.get("B200DP_EXPERIMENTAL", "0") == "0",
                                 reason="experimental kernels: set B200DP_EXPERIMENTAL=1")]

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


@pytest.mark.parametrize("shape", DGRAD_SHAPES)
@pytest.mark.parametrize("cfg", [(0, 1, False), (32, 1, False), (64, 1, False), (128, 1, False), (256, 1, False),
                                 (64, 1, True), (128, 1, True), (256, 1, True), (64, 3, False), (128, 4, True)])
def test_dgrad_with_bn_reduce_matches_the_two_pass_path(shape, cfg):
    """Same dL/da bits as the plain dgrad kernel, and the same sums as bn_bwd_reduce_from_out on it."""
    native = _native()
    from cs744_pytorch_distributed_tutorial_b200.ops._lib import call, ptr, stream_ptr
    B, H, W, C, Cp = shape
    block_n, split_k, pair = cfg
    if block_n and Cp % block_n:
        pytest.skip("block_n does not divide the channel count")
    m_tiles = (B * H * W + 127) // 128
    if pair and m_tiles % 2:
        pytest.skip("the CTA-pair kernel needs an even number of tiles")
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


def test_training_step_with_fused_dgrad_reduction_matches_the_default(monkeypatch):
    """Whole backward with B200DP_DGRAD_BN_REDUCE on vs off: same gradients up to atomics ordering."""
    from cs744_pytorch_distributed_tutorial_b200.models import native_vgg
    from cs744_pytorch_distributed_tutorial_b200.models.vgg import VGG
    torch.manual_seed(0)
    x = torch.randn(64, 3, 32, 32, device=DEV)
    y = torch.randint(0, 10, (64,), device=DEV)
    grads = []
    for flag in (False, True):
        monkeypatch.setattr(native_vgg, "DGRAD_BN_REDUCE", flag)
        torch.manual_seed(1)
        model = VGG("VGG11").to(DEV)
        nv = model.enable_native(max_batch=64)
        nv.forward(x, training=True)
        nv.loss_fwd_bwd(y)
        nv.backward()
        torch.cuda.synchronize()
        grads.append(nv.flat_grads.clone())
    a, b = grads
    cos = torch.nn.functional.cosine_similarity(a, b, dim=0).item()
    assert cos > 0.9995, cos
    assert ((a - b).norm() / b.norm()).item() < 2e-2


def test_part3_weight_relayout_behind_the_buckets(monkeypatch):
    """B200DP_PART3_PREP_OVERLAP=1: after a few captured part3 steps (collective kernels alive on one rank)
    the bf16 operand copies must be exactly the re-layout of the current fp32 weights."""
    monkeypatch.setenv("B200DP_PART3_PREP_OVERLAP", "1")
    from cs744_pytorch_distributed_tutorial_b200.config import TrainConfig
    from cs744_pytorch_distributed_tutorial_b200.engine import Trainer
    cfg = TrainConfig(mode="part3", device="cuda", per_rank_batch=64, eval=False, lr=0.01,
                      single_rank_collectives=True)
    tr = Trainer(cfg)
    try:
        assert tr.backend.sync._prep_overlap and tr.backend.sync.handles_shadows
        g = torch.Generator().manual_seed(5)
        for _ in range(3):
            x = torch.randn(64, 3, 32, 32, generator=g).pin_memory()
            y = torch.randint(0, 10, (64,), generator=g).pin_memory()
            loss = tr.train_step(x, y)
        torch.cuda.synchronize()
        assert torch.isfinite(loss).item()
        nv = tr.backend.native
        got = [(b.wf.clone(), None if b.wd is None else b.wd.clone()) for b in nv.blocks]
        nv.refresh_shadows()
        torch.cuda.synchronize()
        for b, (wf, wd) in zip(nv.blocks, got):
            assert torch.equal(b.wf, wf)
            if wd is not None:
                assert torch.equal(b.wd, wd)
    finally:
        tr.close()


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
    for defer in (False, True):
        torch.manual_seed(1)
        model = VGG("VGG11").to(DEV)
        nv = model.enable_native(max_batch=64)
        nv.defer_head = defer
        for _ in range(2):      # second pass: flags and gradient clearing must have been reset
            nv.forward(x, training=True)
            nv.loss_fwd_bwd(y)
            nv.backward()
        torch.cuda.synchronize()
        res.append((float(nv.loss), int(nv.correct), nv.flat_grads.clone()))
    (l0, c0, g0), (l1, c1, g1) = res
    assert abs(l0 - l1) < 1e-4 and c0 == c1
    assert torch.nn.functional.cosine_similarity(g0, g1, dim=0).item() > 0.9995


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
                      per_rank_batch=32, eval=False, lr=0.01, timeout_s=60)
    tr = Trainer(cfg)
    nv = tr.backend.native
    assert (nv.wgrad_push is not None) == bool(push)
    init = nv.flat_params.cpu()
    g = torch.Generator().manual_seed(300 + rank)
    losses = []
    for _ in range(3):
        x = torch.randn(32, 3, 32, 32, generator=g).pin_memory()
        y = torch.randint(0, 10, (32,), generator=g).pin_memory()
        losses.append(float(tr.train_step(x, y).item()))
    torch.cuda.synchronize()
    torch.save({"params": nv.flat_params.cpu(), "init": init, "losses": losses},
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
    da, db = a["params"] - a["init"], b["params"] - b["init"]
    cos = torch.nn.functional.cosine_similarity(da, db, dim=0).item()
    assert cos > 0.97, cos
    assert abs(db.norm().item() / da.norm().item() - 1) < 0.05


# (B, H, W, Cin, Cout): more tiles than SMs, exactly as many, fewer, and a partly valid last tile
PERSIST_SHAPES = [(256, 16, 16, 64, 128), (64, 8, 8, 128, 256), (296, 4, 4, 256, 512), (32, 4, 4, 512, 512),
                  (30, 2, 2, 512, 512), (8, 32, 32, 64, 64)]


@pytest.mark.parametrize("shape", PERSIST_SHAPES)
@pytest.mark.parametrize("block_n", [64, 128, 256])
@pytest.mark.parametrize("with_stats", [True, False])
def test_persistent_conv_matches_the_one_tile_per_cta_kernel(shape, block_n, with_stats):
    """Same bits as conv_igemm with the same N tile (same accumulation order), same statistics."""
    native = _native()
    B, H, W, Cin, Cout = shape
    if Cout % block_n:
        pytest.skip("block_n does not divide Cout")
    g = torch.Generator(device=DEV).manual_seed(21)
    x = torch.randn(B, H, W, Cin, device=DEV, generator=g).bfloat16()
    w = (torch.randn(Cout, 9 * Cin, device=DEV, generator=g) / (3 * Cin ** 0.5)).bfloat16()
    bias = torch.randn(Cout, device=DEV, generator=g) if with_stats else None
    ref = torch.empty(B, H, W, Cout, device=DEV, dtype=torch.bfloat16)
    st_ref = torch.zeros(2 * Cout, device=DEV) if with_stats else None
    native.conv_igemm(x, w, bias, ref, stats=st_ref, block_n=block_n)
    for rep in range(2):
        out = torch.full_like(ref, float("nan"))
        st = torch.zeros(2 * Cout, device=DEV) if with_stats else None
        native.conv_igemm(x, w, bias, out, stats=st, block_n=block_n, cta_pair=2)
        torch.cuda.synchronize()
        assert torch.equal(out, ref)
        if with_stats:
            torch.testing.assert_close(st, st_ref, rtol=1e-4, atol=1e-3)


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


# (B, H, W, C) of the pooled blocks of VGG-11 at batch 256 / 32, plus ragged batches (partial last stage)
BN_TMA_SHAPES = [(256, 32, 32, 64), (256, 16, 16, 128), (256, 8, 8, 256), (256, 4, 4, 512), (256, 2, 2, 512),
                 (32, 32, 32, 64), (37, 2, 2, 512), (5, 4, 4, 512), (3, 16, 16, 128), (1, 2, 2, 512)]


@pytest.mark.parametrize("shape", BN_TMA_SHAPES)
def test_bulk_copy_pipelined_bn_apply_is_bit_identical(shape):
    """b200dp_bn_bwd_apply_pool_tma vs the register-file kernel it replaces: same dY bits, same dgamma / dbeta,
    and nothing written outside dY (guard bands), twice in a row (barrier phases start over per launch)."""
    from cs744_pytorch_distributed_tutorial_b200.ops._lib import call, ptr, stream_ptr
    B, H, W, C = shape
    g = torch.Generator(device=DEV).manual_seed(31)
    y = torch.randn(B, H, W, C, device=DEV, generator=g).bfloat16()
    dout = torch.randn(B, H // 2, W // 2, C, device=DEV, generator=g).bfloat16()
    mean = torch.randn(C, device=DEV, generator=g) * 0.1
    invstd = torch.rand(C, device=DEV, generator=g) + 0.5
    gamma = torch.rand(C, device=DEV, generator=g) + 0.5
    beta = torch.randn(C, device=DEV, generator=g) * 0.1
    sums = torch.randn(2 * C, device=DEV, generator=g)
    ref = torch.empty_like(y)
    dg0, db0 = torch.empty(C, device=DEV), torch.empty(C, device=DEV)
    call("b200dp_bn_relu_pool_bwd_apply", ptr(y), ptr(dout), ptr(mean), ptr(invstd), ptr(gamma), ptr(beta), ptr(sums),
         ptr(ref), ptr(dg0), ptr(db0), B, H, W, C, 1, stream_ptr())
    n = y.numel()
    for rep in range(2):
        guard = torch.full((n + 4096,), 7.0, device=DEV).bfloat16()
        out = guard[2048: 2048 + n].view(B, H, W, C)
        dg, db = torch.empty(C, device=DEV), torch.empty(C, device=DEV)
        call("b200dp_bn_bwd_apply_pool_tma", ptr(y), ptr(dout), ptr(mean), ptr(invstd), ptr(gamma), ptr(beta), ptr(sums),
             ptr(out), ptr(dg), ptr(db), B, H, W, C, stream_ptr())
        torch.cuda.synchronize()
        assert torch.equal(out, ref)
        assert torch.equal(dg, dg0) and torch.equal(db, db0)
        assert float(guard[:2048].float().min()) == 7.0 and float(guard[2048 + n:].float().max()) == 7.0


def _bn_inputs(B, H, W, C, seed):
    g = torch.Generator(device=DEV).manual_seed(seed)
    y = torch.randn(B, H, W, C, device=DEV, generator=g).bfloat16()
    gamma = torch.rand(C, device=DEV, generator=g) + 0.5
    beta = torch.randn(C, device=DEV, generator=g) * 0.1
    return g, y, gamma, beta


@pytest.mark.parametrize("shape", [(256, 8, 8, 256), (256, 4, 4, 512), (256, 2, 2, 512), (37, 2, 2, 512), (3, 8, 8, 256),
                                   (2, 32, 32, 64)])
def test_bulk_copy_nopool_apply_is_bit_identical(shape):
    from cs744_pytorch_distributed_tutorial_b200.ops._lib import call, ptr, stream_ptr
    B, H, W, C = shape
    g, y, gamma, beta = _bn_inputs(B, H, W, C, 41)
    dout = torch.randn(B, H, W, C, device=DEV, generator=g).bfloat16()
    mean = torch.randn(C, device=DEV, generator=g) * 0.1
    invstd = torch.rand(C, device=DEV, generator=g) + 0.5
    sums = torch.randn(2 * C, device=DEV, generator=g)
    outs = []
    for fn, extra in (("b200dp_bn_relu_pool_bwd_apply", (0,)), ("b200dp_bn_bwd_apply_nopool_tma", ())):
        dy = torch.full_like(y, float("nan"))
        dg, db = torch.empty(C, device=DEV), torch.empty(C, device=DEV)
        for _ in range(2):
            call(fn, ptr(y), ptr(dout), ptr(mean), ptr(invstd), ptr(gamma), ptr(beta), ptr(sums), ptr(dy), ptr(dg),
                 ptr(db), B, H, W, C, *extra, stream_ptr())
        torch.cuda.synchronize()
        outs.append((dy, dg, db))
    for a, b in zip(*outs):
        assert torch.equal(a, b)


@pytest.mark.parametrize("shape", [(256, 32, 32, 64, 1), (256, 16, 16, 128, 1), (256, 8, 8, 256, 0), (256, 8, 8, 256, 1),
                                   (256, 4, 4, 512, 0), (256, 2, 2, 512, 1), (37, 2, 2, 512, 1), (5, 4, 4, 512, 0),
                                   (3, 16, 16, 128, 1)])
@pytest.mark.parametrize("training", [1, 0])
def test_bulk_copy_bn_forward_is_bit_identical(shape, training):
    """Outputs, saved mean / invstd, running statistics and the batch counter, against b200dp_bn_relu_pool_fwd."""
    from cs744_pytorch_distributed_tutorial_b200.ops._lib import call, ptr, stream_ptr
    B, H, W, C, pool = shape
    g, y, gamma, beta = _bn_inputs(B, H, W, C, 43)
    v = y.float().reshape(-1, C)
    stats = torch.cat([v.sum(0), (v * v).sum(0)]).contiguous()
    res = []
    for fn in ("b200dp_bn_relu_pool_fwd", "b200dp_bn_relu_pool_fwd_tma"):
        rm = torch.linspace(-0.2, 0.2, C, device=DEV).contiguous()
        rv = torch.linspace(0.5, 1.5, C, device=DEV).contiguous()
        nbt = torch.full((1,), 5, device=DEV, dtype=torch.int64)
        sm, si = torch.zeros(C, device=DEV), torch.zeros(C, device=DEV)
        ho, wo = (H // 2, W // 2) if pool else (H, W)
        out = torch.full((B, ho, wo, C), float("nan"), device=DEV, dtype=torch.bfloat16)
        call(fn, ptr(y), ptr(stats), ptr(gamma), ptr(beta), ptr(rm), ptr(rv), ptr(nbt), ptr(sm), ptr(si), ptr(out),
             B, H, W, C, pool, training, 0.1, 1e-5, stream_ptr())
        torch.cuda.synchronize()
        res.append((out, rm, rv, nbt, sm, si))
    for a, b in zip(*res):
        assert torch.equal(a, b)
    assert int(res[1][3]) == (6 if training else 5)


@pytest.mark.parametrize("shape", [(256 * 16 * 16, 64), (256 * 8 * 8, 128), (256 * 8 * 8, 256), (256 * 2 * 2, 512),
                                   (256, 512), (37, 512), (1, 64)])
def test_bulk_copy_reduce_from_out_matches(shape):
    from cs744_pytorch_distributed_tutorial_b200.ops._lib import call, ptr, stream_ptr
    Q, C = shape
    g = torch.Generator(device=DEV).manual_seed(47)
    a = torch.relu(torch.randn(Q, C, device=DEV, generator=g)).bfloat16()
    dout = torch.randn(Q, C, device=DEV, generator=g).bfloat16()
    gamma = torch.rand(C, device=DEV, generator=g) + 0.5
    gamma[0] = 0.0
    beta = torch.randn(C, device=DEV, generator=g) * 0.1
    s0, s1 = torch.zeros(2 * C, device=DEV), torch.ones(2 * C, device=DEV)     # the new one accumulates onto ones
    call("b200dp_bn_relu_pool_bwd_reduce_from_out", ptr(a), ptr(dout), ptr(gamma), ptr(beta), ptr(s0), Q, C, stream_ptr())
    call("b200dp_bn_bwd_reduce_from_out_tma", ptr(a), ptr(dout), ptr(gamma), ptr(beta), ptr(s1), Q, C, stream_ptr())
    torch.cuda.synchronize()
    scale = s0.abs().max().item() + 1e-6
    assert ((s1 - 1.0) - s0).abs().max().item() < 1e-4 * scale + 1e-5
    assert abs(s1[C].item() - 1.0) < 1e-6                                        # gamma == 0: no contribution to S2


def _env_worker(rank, world, tag, env, port, outdir):
    import os as _os
    for k, v in env.items():
        _os.environ[k] = v
    _os.environ["LOCAL_RANK"] = str(rank)
    from cs744_pytorch_distributed_tutorial_b200.config import TrainConfig
    from cs744_pytorch_distributed_tutorial_b200.engine import Trainer
    torch.cuda.set_device(rank)
    cfg = TrainConfig(mode="part3", master_ip="127.0.0.1", num_nodes=world, rank=rank, port=port, device="cuda",
                      per_rank_batch=32, eval=False, lr=0.01, timeout_s=60)
    tr = Trainer(cfg)
    nv = tr.backend.native
    init = nv.flat_params.cpu()
    g = torch.Generator().manual_seed(400 + rank)
    losses = []
    for _ in range(4):
        x = torch.randn(32, 3, 32, 32, generator=g).pin_memory()
        y = torch.randint(0, 10, (32,), generator=g).pin_memory()
        losses.append(float(tr.train_step(x, y).item()))
    torch.cuda.synchronize()
    torch.save({"params": nv.flat_params.cpu(), "init": init, "losses": losses}, _os.path.join(outdir, f"{tag}_{rank}.pt"))
    tr.close()


@pytest.mark.multigpu
@pytest.mark.parametrize("env", [{"B200DP_LAZY_EXIT_BARRIER": "1"}, {"B200DP_PART3_PREP_OVERLAP": "1"},
                                 {"B200DP_LAZY_EXIT_BARRIER": "1", "B200DP_WGRAD_PUSH": "1"}])
def test_part3_scheduling_toggles_keep_replicas_identical_and_match_the_default(env):
    import sys
    import tempfile
    import torch.multiprocessing as mp
    sys.path.insert(0, os.path.dirname(__file__))
    from helpers import free_port
    n = torch.cuda.device_count()
    world = 8 if n >= 8 else 4 if n >= 4 else 2
    off = {k: "0" for k in ("B200DP_LAZY_EXIT_BARRIER", "B200DP_PART3_PREP_OVERLAP", "B200DP_WGRAD_PUSH")}
    res = {}
    with tempfile.TemporaryDirectory() as d:
        for tag, e in (("default", off), ("toggled", {**off, **env})):
            mp.spawn(_env_worker, args=(world, tag, e, free_port(), d), nprocs=world, join=True)
            res[tag] = [torch.load(os.path.join(d, f"{tag}_{r}.pt")) for r in range(world)]
    for tag, ranks in res.items():
        for r in ranks[1:]:
            assert torch.equal(ranks[0]["params"], r["params"]), f"{tag}: replicas differ"
    a, b = res["default"][0], res["toggled"][0]
    assert abs(a["losses"][0] - b["losses"][0]) < 1e-2
    da, db = a["params"] - a["init"], b["params"] - b["init"]
    assert torch.nn.functional.cosine_similarity(da, db, dim=0).item() > 0.97
    assert abs(db.norm().item() / da.norm().item() - 1) < 0.05
