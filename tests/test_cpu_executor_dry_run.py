"""Control flow of the GPU executor (models/native_vgg.py) and of the captured training step
(engine.NativeBackend._step_body), walked on CPU.

The kernel wrappers (ops.native) and the CUDA stream / event API are replaced by recorders; tensors live on
the CPU.  What is checked is what cannot be left to a GPU box at the end of a round: that the default step
still issues the kernel sequence the measured build issued, and that each option swaps exactly the calls it
claims to swap -- a NameError or a wrong keyword in `backward()` would otherwise only show up on the hardware.
"""
import contextlib
import types

import pytest
import torch

from cs744_pytorch_distributed_tutorial_b200.models import native_vgg
from cs744_pytorch_distributed_tutorial_b200.models.vgg import VGG


class Recorder:
    """Stands in for ops.native: records (kernel, keyword arguments) and returns its output argument."""

    def __init__(self):
        self.calls = []
        self._lib = types.SimpleNamespace(available=lambda: True)

    def BatchedWeightPrep(self, specs):
        def run():
            self.calls.append(("prep_conv_weights_batched", {"n": len(specs)}))
        return run

    def __getattr__(self, name):
        def kernel(*args, **kw):
            self.calls.append((name, kw))
            return None
        return kernel

    def names(self):
        return [n for n, _ in self.calls]


class FakeEvent:
    def __init__(self, *a, **kw):
        pass

    def record(self, *a):
        pass

    def synchronize(self):
        pass


class FakeStream:
    def __init__(self, *a, **kw):
        pass

    def wait_event(self, ev):
        pass

    def wait_stream(self, s):
        pass


@pytest.fixture
def executor(monkeypatch):
    rec = Recorder()
    monkeypatch.setattr(native_vgg, "K", rec)
    monkeypatch.setattr(native_vgg, "_DRY_RUN", True)
    monkeypatch.setattr(torch.cuda, "Stream", FakeStream)
    monkeypatch.setattr(torch.cuda, "Event", FakeEvent)
    monkeypatch.setattr(torch.cuda, "current_stream", lambda *a: FakeStream())
    monkeypatch.setattr(torch.cuda, "stream", lambda s: contextlib.nullcontext())
    monkeypatch.setattr(torch.cuda, "synchronize", lambda *a: None)

    def make(arch="VGG11", batch=8):
        torch.manual_seed(0)
        model = VGG(arch)
        nv = native_vgg.NativeVGG(model, max_batch=batch, autotune=False)
        rec.calls.clear()
        return model, nv, rec
    return make


def _step(nv, batch=8):
    x = torch.randn(batch, 3, 32, 32)
    y = torch.randint(0, 10, (batch,))
    nv.forward(x, training=True)
    nv.loss_fwd_bwd(y)
    nv.backward()


FWD = ["fill_zero", "im2col3x3_c3"] + ["conv_igemm", "bn_relu_pool_fwd"] * 8 + ["linear_fwd"]
EVAL_FWD = FWD[1:]          # the statistics region is only cleared for a training forward
HEAD = ["cross_entropy", "fill_zero", "linear_bwd"]      # loss, gradient-buffer clear, classifier backward


def test_default_step_issues_the_measured_kernel_sequence(executor):
    model, nv, rec = executor()
    _step(nv)
    names = rec.names()
    assert names[: len(FWD)] == FWD
    assert names[len(FWD): len(FWD) + 3] == HEAD
    bwd = rec.calls[len(FWD) + 3:]
    # per block, deepest first: BN backward, data gradient fused with the BN reduction of the block below (not
    # for block 0), weight gradient
    expect = []
    for i in range(7, -1, -1):
        expect += ["bn_relu_pool_bwd"] + (["conv_dgrad_bn_reduce"] if i > 0 else []) + ["conv_wgrad"]
    assert [n for n, _ in bwd] == expect
    skips = [kw["skip_reduce"] for n, kw in bwd if n == "bn_relu_pool_bwd"]
    assert skips == [False] + [True] * 7                          # only the deepest block reduces on its own
    # a second step behaves the same (no sticky state)
    rec.calls.clear()
    _step(nv)
    assert rec.names()[len(FWD):] == HEAD + expect


def test_eval_forward_and_loss_do_not_touch_the_training_only_kernels(executor):
    model, nv, rec = executor()
    nv.forward(torch.randn(8, 3, 32, 32), training=False)
    nv.loss_fwd_bwd(torch.randint(0, 10, (8,)), need_grad=False)
    assert rec.names() == EVAL_FWD + ["cross_entropy"]


def test_bucket_ready_callbacks_fire_once_per_bucket_in_gradient_order(executor):
    model, nv, rec = executor()
    buckets = nv.layout.buckets("layer")
    nv.set_bucket_plan(buckets)
    fired = []
    nv.on_bucket_ready = fired.append
    _step(nv)
    assert fired == list(range(8))
    nv.set_bucket_plan(nv.layout.buckets("ddp"))
    fired.clear()
    _step(nv)
    assert fired == [0, 1, 2]


def test_fused_head_replaces_exactly_the_three_head_kernels(executor):
    """What the engine's training step runs (it sets ``defer_head``): one head_fwd_bwd launch."""
    model, nv, rec = executor()
    nv.defer_head = True
    _step(nv)
    names = rec.names()
    assert names[: len(FWD) - 1] == FWD[:-1]                      # no linear_fwd at the end of forward
    assert names[len(FWD) - 1: len(FWD) + 1] == ["fill_zero", "head_fwd_bwd"]
    assert "linear_fwd" not in names and "cross_entropy" not in names and "linear_bwd" not in names
    assert names[len(FWD) + 1] == "bn_relu_pool_bwd"
    # an external dlogits (autograd path) still runs the fc backward, and eval still computes logits
    rec.calls.clear()
    nv.forward(torch.randn(8, 3, 32, 32), training=True)
    nv.loss_fwd_bwd(torch.randint(0, 10, (8,)), need_grad=False)
    assert rec.names()[-2:] == ["linear_fwd", "cross_entropy"]
    rec.calls.clear()
    nv.forward(torch.randn(8, 3, 32, 32), training=False)
    assert rec.names()[-1] == "linear_fwd"


def test_wgrad_push_configuration_reroutes_the_configured_blocks_only(executor):
    model, nv, rec = executor()
    nv.wgrad_push = {i: {"marker": i} for i in (3, 5)}
    _step(nv)
    pushed = [kw["marker"] for n, kw in rec.calls if n == "conv_wgrad_push"]
    assert pushed == [5, 3]
    assert rec.names().count("conv_wgrad") == 6


@pytest.mark.parametrize("arch,blocks", [("VGG13", 10), ("VGG16", 13), ("VGG19", 16)])
def test_deeper_architectures_walk_the_same_executor(executor, arch, blocks):
    model, nv, rec = executor(arch, batch=4)
    _step(nv, batch=4)
    names = rec.names()
    assert names.count("bn_relu_pool_fwd") == blocks and names.count("conv_wgrad") == blocks
    assert names.count("conv_igemm") == blocks and names.count("conv_dgrad_bn_reduce") == blocks - 1


def test_engine_step_body_runs_end_to_end_in_dry_run(executor, monkeypatch):
    """engine.NativeBackend at world size 1 (the configuration of smoke() and of the 1-GPU bench): constructor,
    LocalSGD two-phase optimizer, staged inputs, one eager step -- every Python line the round-end GPU run
    executes before and between kernel launches."""
    from cs744_pytorch_distributed_tutorial_b200 import engine
    from cs744_pytorch_distributed_tutorial_b200.config import TrainConfig
    from cs744_pytorch_distributed_tutorial_b200.parallel import sync_native
    model, nv0, rec = executor()          # installs the recorders; this executor instance itself is not used
    monkeypatch.setattr(sync_native, "K", rec)
    monkeypatch.setattr(engine, "K", rec)
    monkeypatch.setattr(torch.cuda, "set_device", lambda d: None)
    real_zeros = torch.zeros
    monkeypatch.setattr(torch, "zeros", lambda *a, **kw: real_zeros(*a, **{k: v for k, v in kw.items() if k != "pin_memory"}))
    cfg = TrainConfig(mode="part3", device="cpu", per_rank_batch=8, eval=False, cuda_graph=False)
    orig_enable = VGG.enable_native
    monkeypatch.setattr(VGG, "enable_native", lambda self, max_batch, heap=None, **kw: orig_enable(self, max_batch, heap=heap, autotune=False, **kw))
    be = engine.NativeBackend(cfg, VGG("VGG11"), torch.device("cpu"))
    assert type(be.sync).__name__ == "LocalSGD"
    rec.calls.clear()
    loss = be.train_step(torch.randn(8, 3, 32, 32), torch.randint(0, 10, (8,)))
    names = rec.names()
    assert loss is be.native.loss and be.steps_done == 1
    assert names[0] == "copy_pair"                         # staged batch -> the step's static input buffers
    assert names[1: len(FWD)] == FWD[:-1] and names.count("head_fwd_bwd") == 1 and names.count("conv_wgrad") == 8
    assert names.count("sgd_momentum") == 2 and names.count("prep_conv_weights_batched") == 2   # two-phase optimizer
    assert be.input_event() is not None
    # ragged last batch: same kernels, eager
    rec.calls.clear()
    be.train_step(torch.randn(5, 3, 32, 32), torch.randint(0, 10, (5,)))
    assert rec.names().count("conv_wgrad") == 8 and be.steps_done == 2


def test_engine_part3_two_rank_path_in_dry_run(executor, monkeypatch):
    """The world > 1 part3 configuration (what the scaling bench runs): store + communicator + bucket reducer
    replaced by stand-ins, one step walked on CPU.  Checks the order of the collective-side calls around the
    compute kernels: parameter / BatchNorm broadcast at start-up, BatchNorm-state broadcast after every
    forward, one bucket_ready per bucket during backward, finalize, weight re-layout."""
    from cs744_pytorch_distributed_tutorial_b200 import engine
    from cs744_pytorch_distributed_tutorial_b200.config import TrainConfig
    from cs744_pytorch_distributed_tutorial_b200.parallel import comm as comm_mod
    from cs744_pytorch_distributed_tutorial_b200.parallel import sync_native
    model, nv0, rec = executor()
    log = rec.calls

    class FakeStore:
        def __init__(self, *a, **kw):
            pass

        def barrier(self, name="b"):
            log.append(("store.barrier", {"name": name}))

        def close(self):
            pass

    class FakeCommunicator:
        def __init__(self, store, rank, world, device, heap_bytes, timeout_s=20.0, multicast=False):
            self.rank, self.world, self.multicast, self.heap_bytes = rank, world, multicast, heap_bytes
            self.has_multicast = multicast

        def prefer_nvls(self, on=True):
            log.append(("comm.prefer_nvls", {"on": on}))

        def zeros(self, numel, dtype=torch.float32, align=256):
            return torch.zeros(numel, dtype=dtype)

        def offset_of(self, t):
            return 0

        def broadcast(self, t, root=0, blocks=0, channel=0, stream=None):
            log.append(("comm.broadcast", {"numel": t.numel(), "channel": channel}))

        def close(self):
            pass

    class FakeReducer:
        def __init__(self, comm, buckets, *a, **kw):
            self.n_buckets, self.kw = len(buckets), kw

        def bucket_ready(self, k):
            log.append(("reducer.bucket_ready", {"k": k}))

        def finalize(self):
            log.append(("reducer.finalize", {}))

        def set_lazy_exit(self, on=True):
            log.append(("reducer.set_lazy_exit", {"on": on}))

        def stream(self, device):
            return FakeStream()

        def close(self):
            pass

    monkeypatch.setattr(comm_mod, "Store", FakeStore)
    monkeypatch.setattr(comm_mod, "Communicator", FakeCommunicator)
    monkeypatch.setattr(sync_native, "BucketReducer", FakeReducer)
    monkeypatch.setattr(sync_native, "K", rec)
    monkeypatch.setattr(engine, "K", rec)
    monkeypatch.setattr(torch.cuda, "set_device", lambda d: None)
    real_zeros = torch.zeros
    monkeypatch.setattr(torch, "zeros", lambda *a, **kw: real_zeros(*a, **{k: v for k, v in kw.items() if k != "pin_memory"}))
    orig_enable = VGG.enable_native
    monkeypatch.setattr(VGG, "enable_native", lambda self, max_batch, heap=None, **kw: orig_enable(self, max_batch, heap=heap, autotune=False, **kw))
    monkeypatch.delenv("B200DP_WGRAD_PUSH", raising=False)

    cfg = TrainConfig(mode="part3", device="cpu", num_nodes=2, rank=1, per_rank_batch=8, eval=False, cuda_graph=False,
                      bucket_plan="layer")
    log.clear()
    be = engine.NativeBackend(cfg, VGG("VGG11"), torch.device("cpu"))
    assert type(be.sync).__name__ == "BucketedOverlapSync" and be.comm.multicast      # NVLS heap by default for world > 1
    assert be.sync.reducer.kw["blocks"] == 32
    start = [n for n, _ in log]
    assert start.count("comm.broadcast") == 2 and "comm.prefer_nvls" in start         # parameters + BatchNorm state
    log.clear()
    be.train_step(torch.randn(8, 3, 32, 32), torch.randint(0, 10, (8,)))
    names = [n for n, _ in log]
    i_fwd_end = max(i for i, n in enumerate(names) if n == "bn_relu_pool_fwd")
    assert names[i_fwd_end + 1] == "comm.broadcast"                                   # BatchNorm state, after forward
    ready = [kw["k"] for n, kw in log if n == "reducer.bucket_ready"]
    assert ready == list(range(8))
    i_fin = names.index("reducer.finalize")
    assert i_fin > max(i for i, n in enumerate(names) if n == "conv_wgrad")
    assert "prep_conv_weights_batched" in names[i_fin:]                               # re-layout after the fused SGD
    assert "sgd_momentum" not in names                                                # the optimizer lives in the collectives
    be.close()
