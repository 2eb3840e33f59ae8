"""part3 across W "ranks" in one process: the real executors and the real BucketedOverlapSync strategy, with
the kernels replaced by fp32 torch restatements (tests/emulated_kernels.py) and the NVLink side by an
in-process fabric (tests/multirank_fabric.py).

What this pins down without a multi-GPU box: that the strategy hands the collectives the right ranges, that
replicas stay identical, that the result is DistributedDataParallel's (mean of the per-rank gradients, SGD with
the reference's hyper-parameters; ref: master/part3/part3.py:116-117), and -- the reason this file exists --
that the experimental weight-gradient -> reduce-scatter path (B200DP_WGRAD_PUSH) produces the same parameters
as the bucket all-reduce, with every staged float4 written by exactly the push that should write it.
"""
import contextlib
import copy
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.dirname(__file__))
from emulated_kernels import EmulatedKernels  # noqa: E402
from multirank_fabric import Fabric, FabricComm, FabricReducer, push_tile_handoff  # noqa: E402

from cs744_pytorch_distributed_tutorial_b200.models import native_vgg  # noqa: E402
from cs744_pytorch_distributed_tutorial_b200.models.vgg import VGG  # noqa: E402
from cs744_pytorch_distributed_tutorial_b200.parallel import sync_native  # noqa: E402

LR, MOM, WD = 0.05, 0.9, 1e-4


class FabricKernels(EmulatedKernels):
    fabric = None

    def conv_wgrad_push(self, dy, x, dw, *, counters, world, rank, peer_stage, flat_off, own_start, own_len, stage_base,
                        split_k=0, block_n=0):
        self.conv_wgrad(dy, x, dw)
        self.calls[-1] = "conv_wgrad_push"
        push_tile_handoff(self.fabric, dw.reshape(-1), world=world, rank=rank, peer_stage=peer_stage, flat_off=flat_off,
                          own_start=own_start, own_len=own_len, stage_base=stage_base)
        return dw


class _Ev:
    def __init__(self, *a, **kw):
        pass

    def record(self, *a):
        pass

    def synchronize(self):
        pass


class _St:
    def __init__(self, *a, **kw):
        pass

    def wait_event(self, e):
        pass

    def wait_stream(self, s):
        pass


@pytest.fixture
def emu(monkeypatch):
    k = FabricKernels()
    for mod in (native_vgg, sync_native):
        monkeypatch.setattr(mod, "K", k)
    monkeypatch.setattr(native_vgg, "_DRY_RUN", True)
    monkeypatch.setattr(native_vgg, "_ACT_DTYPE", torch.float32)
    monkeypatch.setattr(sync_native, "BucketReducer", FabricReducer)
    monkeypatch.setattr(torch.cuda, "Stream", _St)
    monkeypatch.setattr(torch.cuda, "Event", _Ev)
    monkeypatch.setattr(torch.cuda, "current_stream", lambda *a: _St())
    monkeypatch.setattr(torch.cuda, "stream", lambda s: contextlib.nullcontext())
    monkeypatch.setattr(torch.cuda, "synchronize", lambda *a: None)
    monkeypatch.delenv("B200DP_WGRAD_PUSH", raising=False)
    return k


def _base_model(seed=0):
    torch.manual_seed(seed)
    m = VGG("VGG11")
    with torch.no_grad():
        for mod in m.modules():
            if isinstance(mod, torch.nn.BatchNorm2d):
                mod.weight.uniform_(0.5, 1.5)
                mod.bias.uniform_(-0.3, 0.3)
    return m


def _run(emu, world, batches, bucket_plan="layer", mode="part3", fuse=True):
    """`batches`: list over steps of list over ranks of (x, y).  Returns the fabric (executors, strategies)."""
    fabric = Fabric(world)
    emu.fabric = fabric
    base = _base_model()
    B = batches[0][0][0].shape[0]
    for r in range(world):
        model = copy.deepcopy(base)
        comm = FabricComm(fabric, r)
        nv = native_vgg.NativeVGG(model, max_batch=B, heap=comm, autotune=False)
        fabric.execs[r] = nv
        kw = dict(lr=LR, momentum=MOM, weight_decay=WD, fuse_optimizer=fuse)
        if mode == "part3":
            sync = sync_native.BucketedOverlapSync(nv, comm, bucket_plan=bucket_plan, blocks=32, **kw)
        elif mode == "part2b":
            sync = sync_native.AllReduceSync(nv, comm, **kw)
        else:
            sync = sync_native.ParamServerSync(nv, comm, push=(mode == "part2a_extra"), **kw)
        fabric.syncs[r] = sync
    for r in range(world):
        fabric.syncs[r].on_init()
    for step in batches:
        for r in range(world):
            s = fabric.syncs[r]
            if getattr(s, "_wgrad_push", False):
                s._stage.fill_(float("nan"))          # a slot nobody writes must show up in the parameters
        for r, (x, y) in enumerate(step):
            nv, s = fabric.execs[r], fabric.syncs[r]
            s.pre_forward()
            nv.forward(x, training=True)
            s.post_forward()
            nv.loss_fwd_bwd(y)
            nv.backward()
        if mode == "part3":
            fabric.run_collectives()
        for r in range(world):
            nv, s = fabric.execs[r], fabric.syncs[r]
            s.sync_and_step()
            if not s.handles_shadows:
                nv.refresh_shadows()
            s.pre_forward()
    return fabric


def _reference(world, batches):
    """DistributedDataParallel semantics on one torch module: per-rank forward/backward (per-rank BatchNorm
    statistics), gradients averaged over the ranks, torch.optim.SGD."""
    ref = _base_model().train()
    opt = torch.optim.SGD(ref.parameters(), lr=LR, momentum=MOM, weight_decay=WD)
    for step in batches:
        acc = [torch.zeros_like(p) for p in ref.parameters()]
        for x, y in step:
            ref.zero_grad()
            torch.nn.functional.cross_entropy(ref(x), y).backward()
            for a, p in zip(acc, ref.parameters()):
                a += p.grad / world
        for a, p in zip(acc, ref.parameters()):
            p.grad = a
        opt.step()
    return ref


def _batches(world, steps, B=4, seed=3):
    g = torch.Generator().manual_seed(seed)
    return [[(torch.randn(B, 3, 32, 32, generator=g), torch.randint(0, 10, (B,), generator=g)) for _ in range(world)]
            for _ in range(steps)]


def _check_against_reference(fabric, ref):
    nv0 = fabric.execs[0]
    for nv in fabric.execs[1:]:
        assert torch.equal(nv.flat_params, nv0.flat_params), "replicas differ"
    assert torch.isfinite(nv0.flat_params).all()
    ref_params = dict(ref.named_parameters())
    for name, p in nv0.module.named_parameters():
        a, b = p.data.flatten(), ref_params[name].data.flatten()
        rel = ((a - b).norm() / (b.norm() + 1e-12)).item()
        assert rel < 1e-2, f"{name}: relative L2 {rel:.3e} against the DDP reference"


@pytest.mark.parametrize("world", [2, 3])
def test_bucketed_allreduce_path_is_ddp(emu, world):
    batches = _batches(world, steps=2)
    fabric = _run(emu, world, batches)
    _check_against_reference(fabric, _reference(world, batches))
    # DDP semantics for the buffers: every rank starts the next forward from rank 0's BatchNorm state
    for nv in fabric.execs[1:]:
        assert torch.equal(nv.bn_state, fabric.execs[0].bn_state)


@pytest.mark.parametrize("world", [2, 3, 4])
def test_wgrad_push_path_gives_the_same_parameters_as_the_allreduce(emu, monkeypatch, world):
    batches = _batches(world, steps=2, B=2)
    plain = _run(emu, world, batches)
    want = plain.execs[0].flat_params.clone()
    monkeypatch.setenv("B200DP_WGRAD_PUSH", "1")
    pushed = _run(emu, world, batches)
    assert all(s._wgrad_push for s in pushed.syncs)
    assert emu.calls.count("conv_wgrad_push") > 0
    got = pushed.execs[0].flat_params
    for nv in pushed.execs[1:]:
        assert torch.equal(nv.flat_params, got), "replicas differ"
    assert torch.isfinite(got).all(), "a staged float4 was read that no rank had written"
    assert (got - want).abs().max().item() <= 1e-5 * want.abs().max().item()
    _check_against_reference(pushed, _reference(world, batches))


@pytest.mark.parametrize("mode,fuse", [("part2a", True), ("part2a_extra", True), ("part2b", True), ("part2b", False),
                                       ("part3", False)])
def test_every_synchronisation_mode_reaches_the_same_parameters(emu, mode, fuse):
    """The tutorial's point: gather/scatter, point-to-point parameter server, all_reduce and DDP are the same
    optimisation step (ref: master/part2a/part2a.py:42-53, part2a_extra.py:41-58, part2b/part2b.py:43-46,
    part3/part3.py:116-117)."""
    world = 3
    batches = _batches(world, steps=2, B=2)
    fabric = _run(emu, world, batches, mode=mode, fuse=fuse)
    _check_against_reference(fabric, _reference(world, batches))


@pytest.mark.parametrize("mode", ["part2b", "part3"])
def test_sharded_momentum_is_gathered_for_the_checkpoint(emu, mode):
    """The fused collectives keep SGD momentum only on the owner of each 1/W slice.  A checkpoint must hold
    what torch.optim.SGD holds (ref: optimizer at master/part3/part3.py:117-118), so every rank gathers the
    owners' slices first; before the gather a non-owner's copy is stale."""
    world = 3
    batches = _batches(world, steps=2, B=2)
    fabric = _run(emu, world, batches, mode=mode, fuse=True)
    nv0, nv1 = fabric.execs[0], fabric.execs[1]
    assert not torch.equal(nv0.flat_momentum, nv1.flat_momentum), "the emulation no longer models slice ownership"
    for s in fabric.syncs:
        s.gather_optimizer_state()
    for nv in fabric.execs[1:]:
        assert torch.equal(nv.flat_momentum, nv0.flat_momentum)
    # against torch.optim.SGD's buffers after the same two steps
    ref = _base_model().train()
    opt = torch.optim.SGD(ref.parameters(), lr=LR, momentum=MOM, weight_decay=WD)
    for step in batches:
        acc = [torch.zeros_like(p) for p in ref.parameters()]
        for x, y in step:
            ref.zero_grad()
            torch.nn.functional.cross_entropy(ref(x), y).backward()
            for a, p in zip(acc, ref.parameters()):
                a += p.grad / world
        for a, p in zip(acc, ref.parameters()):
            p.grad = a
        opt.step()
    ref_named = dict(ref.named_parameters())
    for e in nv0.layout.entries:
        got = nv0._view(nv0.flat_momentum, e).flatten()
        want = opt.state[ref_named[e.name]]["momentum_buffer"].flatten()
        rel = ((got - want).norm() / (want.norm() + 1e-12)).item()
        assert rel < 2e-2, f"momentum of {e.name}: relative L2 {rel:.3e} against torch.optim.SGD"
