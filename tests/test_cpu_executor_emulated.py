"""The GPU executor's data flow, run numerically on the host.

models/native_vgg.py decides which buffer goes into which kernel; the kernels themselves are validated on the
GPU against torch, one by one.  Here the kernel wrappers are replaced by fp32 torch restatements with the same
in-place contracts (tests/emulated_kernels.py) and the whole forward / loss / backward of the executor is
compared with torch autograd on the reference module -- for the default flow and for every experimental
toggle.  A wrong buffer, a missing clear, a reduction fed from the wrong activation shows up as a gradient
mismatch without needing a GPU.
"""
import contextlib
import copy
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.dirname(__file__))
from emulated_kernels import EmulatedKernels  # noqa: E402

from cs744_pytorch_distributed_tutorial_b200.models import native_vgg  # noqa: E402
from cs744_pytorch_distributed_tutorial_b200.models.vgg import VGG  # noqa: E402


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
    k = EmulatedKernels()
    from cs744_pytorch_distributed_tutorial_b200 import engine
    monkeypatch.setattr(native_vgg, "K", k)
    monkeypatch.setattr(engine, "K", k)
    monkeypatch.setattr(native_vgg, "_DRY_RUN", True)
    monkeypatch.setattr(native_vgg, "_ACT_DTYPE", torch.float32)
    monkeypatch.setattr(torch.cuda, "Stream", _St)
    monkeypatch.setattr(torch.cuda, "Event", _Ev)
    monkeypatch.setattr(torch.cuda, "current_stream", lambda *a: _St())
    monkeypatch.setattr(torch.cuda, "stream", lambda s: contextlib.nullcontext())
    monkeypatch.setattr(torch.cuda, "synchronize", lambda *a: None)
    return k


def _reference(model, x, y):
    ref = copy.deepcopy(model).train()
    out = ref(x)
    loss = torch.nn.functional.cross_entropy(out, y)
    loss.backward()
    return ref, out.detach(), loss.detach()


def _same_as_default(nv, x, y, monkeypatch, tol=2e-5):
    """Re-run the step on the plain flow (three-kernel head, no push) and require the same gradients: the
    options re-route kernels, they must not change the maths beyond summation order."""
    got = nv.flat_grads.clone()
    loss = float(nv.loss)
    nv.defer_head, nv.wgrad_push = False, None
    _step(nv, x, y)
    scale = nv.flat_grads.abs().max().item()
    assert (got - nv.flat_grads).abs().max().item() <= tol * scale
    assert abs(loss - float(nv.loss)) <= 1e-6


def _build(arch="VGG11", batch=6, seed=0):
    torch.manual_seed(seed)
    model = VGG(arch)
    with torch.no_grad():                     # non-trivial BatchNorm affine so that gamma / beta matter
        for m in model.modules():
            if isinstance(m, torch.nn.BatchNorm2d):
                m.weight.uniform_(0.5, 1.5)
                m.bias.uniform_(-0.3, 0.3)
    x = torch.randn(batch, 3, 32, 32)
    y = torch.randint(0, 10, (batch,))
    ref, ref_logits, ref_loss = _reference(model, x, y)
    nv = native_vgg.NativeVGG(model, max_batch=batch, autotune=False)
    return model, nv, ref, ref_logits, ref_loss, x, y


def _compare(model, nv, ref, ref_logits, ref_loss, tol=3e-3, buffers=True):
    """Against torch autograd.  Per tensor: relative L2 error and cosine.  A pre-activation within rounding of
    zero can land on the other side of the ReLU in the two implementations (about one element per step at
    these sizes), which moves a few gradient entries by O(1/P); a wiring mistake moves everything by O(1)."""
    torch.testing.assert_close(nv.logits[: ref_logits.shape[0]], ref_logits, rtol=tol, atol=tol)
    torch.testing.assert_close(nv.loss, ref_loss, rtol=tol, atol=tol)
    ref_params = dict(ref.named_parameters())
    for name, p in model.named_parameters():
        g, gr = p.grad.flatten(), ref_params[name].grad.flatten()
        if gr.norm().item() < 1e-5:                 # conv biases in front of a training-mode BatchNorm: exactly zero
            assert g.norm().item() < 1e-4, name
            continue
        rel = ((g - gr).norm() / gr.norm()).item()
        cos = torch.nn.functional.cosine_similarity(g, gr, dim=0).item()
        assert rel < 3e-2 and cos > 0.9995, f"{name}: relative L2 error {rel:.3e}, cosine {cos:.6f}"
    for (n, b), (_, br) in zip(model.named_buffers(), ref.named_buffers()):
        if buffers:
            torch.testing.assert_close(b.float(), br.float(), rtol=1e-4, atol=1e-5, msg=f"buffer {n}")


def _step(nv, x, y):
    nv.forward(x, training=True)
    nv.loss_fwd_bwd(y)
    nv.backward()


def test_default_flow_matches_autograd(emu):
    """Including the data gradient that also reduces the BatchNorm backward of the block below."""
    model, nv, ref, ref_logits, ref_loss, x, y = _build()
    _step(nv, x, y)
    assert emu.calls.count("conv_dgrad_bn_reduce") == 7
    _compare(model, nv, ref, ref_logits, ref_loss)
    assert nv.grad_views_intact()


def test_fused_head_flow_matches_autograd(emu, monkeypatch):
    model, nv, ref, ref_logits, ref_loss, x, y = _build(seed=2)
    nv.defer_head = True
    _step(nv, x, y)
    assert "head_fwd_bwd" in emu.calls and "linear_bwd" not in emu.calls
    _compare(model, nv, ref, ref_logits, ref_loss)
    assert int(nv.correct) == int((ref_logits.argmax(1) == y).sum())
    _same_as_default(nv, x, y, monkeypatch)


def test_all_options_together_and_two_steps(emu, monkeypatch):
    """Second step after a parameter change: buffers that must be cleared per step (statistics, sums,
    gradients) really are, with the fused head and the push flavour of the weight gradient switched on."""
    model, nv, ref, ref_logits, ref_loss, x, y = _build(seed=3)
    nv.defer_head = True
    nv.wgrad_push = {i: dict(counters=None, world=2, rank=0, peer_stage=None, flat_off=0, own_start=0, own_len=0,
                             stage_base=0) for i in range(1, 8)}
    _step(nv, x, y)
    _compare(model, nv, ref, ref_logits, ref_loss)
    # same perturbation on both sides, fresh batch
    torch.manual_seed(9)
    with torch.no_grad():
        ref_params = dict(ref.named_parameters())
        for name, p in model.named_parameters():
            d = 0.01 * torch.randn_like(ref_params[name])
            ref_params[name].add_(d)
            p.data.add_(d)
    nv.refresh_shadows()
    x2, y2 = torch.randn_like(x), torch.randint(0, 10, y.shape)
    ref.zero_grad()
    out = ref(x2)
    loss = torch.nn.functional.cross_entropy(out, y2)
    loss.backward()
    _step(nv, x2, y2)
    _compare(model, nv, ref, out.detach(), loss.detach())
    _same_as_default(nv, x2, y2, monkeypatch)


def test_eval_forward_uses_the_running_statistics(emu):
    model, nv, ref, ref_logits, ref_loss, x, y = _build(seed=4)
    _step(nv, x, y)                      # one training step moves the running statistics on both sides
    ref.eval()
    with torch.no_grad():
        want = ref(x)
    got = nv.forward(x, training=False)
    torch.testing.assert_close(got, want, rtol=2e-4, atol=2e-4)


def test_state_dict_round_trip_through_the_flat_buffers(emu):
    model, nv, ref, _, _, x, y = _build(seed=5)
    _step(nv, x, y)                      # the reference already ran this step: same running statistics now
    sd = model.portable_state_dict()
    ref_sd = ref.state_dict()
    assert list(sd) == list(ref_sd) and len(sd) == 58
    for k in sd:
        torch.testing.assert_close(sd[k].float(), ref_sd[k].float(), msg=k)      # OIHW view of KRSC memory
    other = VGG("VGG11")
    other.load_state_dict(sd)
    x = torch.randn(2, 3, 32, 32)
    other.eval(), ref.eval()
    with torch.no_grad():
        torch.testing.assert_close(other(x), ref(x))


def test_deeper_architecture_and_bf16_activations(emu, monkeypatch):
    """VGG16 through the same executor; with bf16 activation buffers the gradients stay aligned with fp32
    autograd (cosine), the precision the real kernels work at."""
    model, nv, ref, ref_logits, ref_loss, x, y = _build("VGG16", batch=4, seed=6)
    _step(nv, x, y)
    _compare(model, nv, ref, ref_logits, ref_loss)
    monkeypatch.setattr(native_vgg, "_ACT_DTYPE", torch.bfloat16)
    model, nv, ref, ref_logits, ref_loss, x, y = _build("VGG11", batch=8, seed=7)
    _step(nv, x, y)
    ref_params = dict(ref.named_parameters())
    a = torch.cat([p.grad.flatten() for n, p in model.named_parameters() if "weight" in n and p.dim() == 4])
    b = torch.cat([ref_params[n].grad.flatten() for n, p in model.named_parameters() if "weight" in n and p.dim() == 4])
    assert torch.nn.functional.cosine_similarity(a, b, dim=0).item() > 0.9
    assert abs(float(nv.loss) - float(ref_loss)) < 0.05


def test_engine_training_steps_match_torch_sgd(emu, monkeypatch):
    """engine.NativeBackend (one rank: two-phase fused SGD + operand re-layout) for three steps against
    torch.optim.SGD with the reference's hyper-parameters (ref: master/part1/part1.py:98-99): parameters,
    momentum behaviour and BatchNorm state stay together."""
    from cs744_pytorch_distributed_tutorial_b200 import engine
    from cs744_pytorch_distributed_tutorial_b200.config import TrainConfig
    from cs744_pytorch_distributed_tutorial_b200.parallel import sync_native
    monkeypatch.setattr(sync_native, "K", emu)
    monkeypatch.setattr(torch.cuda, "set_device", lambda d: None)
    real_zeros = torch.zeros
    monkeypatch.setattr(torch, "zeros", lambda *a, **kw: real_zeros(*a, **{k: v for k, v in kw.items() if k != "pin_memory"}))
    orig_enable = VGG.enable_native
    monkeypatch.setattr(VGG, "enable_native", lambda self, max_batch, heap=None, **kw: orig_enable(self, max_batch, heap=heap, autotune=False, **kw))
    torch.manual_seed(11)
    model = VGG("VGG11")
    ref = copy.deepcopy(model).train()
    opt = torch.optim.SGD(ref.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-4)
    cfg = TrainConfig(mode="part1", device="cpu", per_rank_batch=8, eval=False, cuda_graph=False, lr=0.05)
    be = engine.NativeBackend(cfg, model, torch.device("cpu"))
    for step in range(3):
        x, y = torch.randn(8, 3, 32, 32), torch.randint(0, 10, (8,))
        opt.zero_grad()
        loss_ref = torch.nn.functional.cross_entropy(ref(x), y)
        loss_ref.backward()
        opt.step()
        loss = be.train_step(x, y)
        assert abs(float(loss) - float(loss_ref)) < 5e-3 * max(1.0, abs(float(loss_ref))), (step, float(loss), float(loss_ref))
    ref_params = dict(ref.named_parameters())
    for name, p in model.named_parameters():
        a, b = p.data.flatten(), ref_params[name].data.flatten()
        rel = ((a - b).norm() / (b.norm() + 1e-12)).item()
        assert rel < 2e-2, f"{name}: parameters drifted apart, relative L2 {rel:.3e}"
    # the bf16-operand copies the next forward will read are the re-layout of the CURRENT weights
    nv = be.native
    got = [(b.wf.clone(), None if b.wd is None else b.wd.clone()) for b in nv.blocks]
    nv.refresh_shadows()
    for b, (wf, wd) in zip(nv.blocks, got):
        assert torch.equal(b.wf, wf) and (wd is None or torch.equal(b.wd, wd))


def test_trainer_fit_evaluate_checkpoint_resume_on_the_native_backend(emu, monkeypatch, tmp_path, capsys):
    """The tutorial's whole job through Trainer on the native backend (emulated kernels): a ragged last batch,
    the test pass, a checkpoint in the reference's state_dict format, and a resumed trainer that continues
    from the same parameters and momentum (ref: master/part1/part1.py:65-103)."""
    from cs744_pytorch_distributed_tutorial_b200 import engine
    from cs744_pytorch_distributed_tutorial_b200.config import TrainConfig
    from cs744_pytorch_distributed_tutorial_b200.parallel import sync_native
    monkeypatch.setattr(sync_native, "K", emu)
    monkeypatch.setattr(torch.cuda, "set_device", lambda d: None)
    real_zeros = torch.zeros
    monkeypatch.setattr(torch, "zeros", lambda *a, **kw: real_zeros(*a, **{k: v for k, v in kw.items() if k != "pin_memory"}))
    orig_enable = VGG.enable_native
    monkeypatch.setattr(VGG, "enable_native", lambda self, max_batch, heap=None, **kw: orig_enable(self, max_batch, heap=heap, autotune=False, **kw))
    ck = str(tmp_path / "ck.pt")
    for loader in ("python", "native"):
        cfg = TrainConfig(mode="part1", device="cpu", backend="native", global_batch=8, train_size=20, test_size=12,
                          lr=0.01, cuda_graph=False, checkpoint=ck, loader=loader, loader_threads=2, json_metrics=False)
        tr = engine.Trainer(cfg)
        out = tr.fit()
        assert tr.backend_name == "native" and out["iters"] == 3 and out["samples"] == 12    # 8 + 8 + 4 images
        assert tr.backend.steps_done == 3 and 0 <= out["correct"] <= 12
        params = tr.backend.native.flat_params.clone()
        mom = tr.backend.native.flat_momentum.clone()
        tr.close()
        lines = capsys.readouterr().out.splitlines()
        assert lines[0].startswith("0 loss: ") and any(l.startswith("Test set: Average loss:") for l in lines)
        # reference-format checkpoint: plain torch module loads it
        payload = torch.load(ck, weights_only=True)
        plain = VGG("VGG11")
        plain.load_state_dict(payload["model"])
        assert len(payload["model"]) == 58 and payload["step"] == 3
        # resume: same weights, same momentum, shadows refreshed
        cfg2 = TrainConfig(mode="part1", device="cpu", backend="native", global_batch=8, train_size=20, test_size=12,
                           lr=0.01, cuda_graph=False, resume=ck, json_metrics=False)
        tr2 = engine.Trainer(cfg2)
        nv2 = tr2.backend.native
        torch.testing.assert_close(nv2.flat_params[: nv2.layout.used], params[: nv2.layout.used])
        torch.testing.assert_close(nv2.flat_momentum[: nv2.layout.used], mom[: nv2.layout.used])
        assert tr2.global_step == 3
        got = [b.wf.clone() for b in nv2.blocks]
        nv2.refresh_shadows()
        assert all(torch.equal(b.wf, w) for b, w in zip(nv2.blocks, got))
        tr2.close()


def test_autograd_bridge_drives_the_native_backward(emu):
    """The eager user API: logits = model(x); loss.backward() (torch autograd calling into the executor through
    _NativeVGGFunction), then the fused optimizer wrapper -- gradients land in the flat buffer views that the
    module's parameters already point at, once, not doubled."""
    model, nv, ref, ref_logits, ref_loss, x, y = _build(seed=13)
    logits = nv.autograd_forward(x, True)
    loss = torch.nn.functional.cross_entropy(logits, y)
    loss.backward()
    torch.testing.assert_close(logits.detach(), ref_logits, rtol=3e-3, atol=3e-3)
    assert nv.grad_views_intact()
    ref_params = dict(ref.named_parameters())
    for name, p in model.named_parameters():
        g, gr = p.grad.flatten(), ref_params[name].grad.flatten()
        if gr.norm().item() < 1e-5:
            continue
        rel = ((g - gr).norm() / gr.norm()).item()
        assert rel < 3e-2, f"{name}: relative L2 error {rel:.3e} (doubled gradients would give 1.0)"
    # a second backward overwrites (the tutorial zeroes gradients every iteration), it does not accumulate
    first = nv.flat_grads.clone()
    logits = nv.autograd_forward(x, True)
    torch.nn.functional.cross_entropy(logits, y).backward()
    assert (nv.flat_grads - first).abs().max().item() <= 1e-5 * first.abs().max().item()


def test_eager_training_loop_with_fused_sgd_matches_torch(emu, monkeypatch):
    """The tutorial's loop verbatim -- zero_grad / forward / loss / backward / step (ref: master/part1/part1.py:
    31-38) -- on a native-enabled model with ops.optim.FusedSGD, against the same loop on the plain module
    with torch.optim.SGD, including a foreign zero_grad(set_to_none=True) that drops the gradient views."""
    from cs744_pytorch_distributed_tutorial_b200.ops import optim as fused_optim
    monkeypatch.setattr(fused_optim, "K", emu)
    model, nv, ref, _, _, x, y = _build(seed=14)
    model._native = nv
    ref.zero_grad()
    opt_ref = torch.optim.SGD(ref.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-4)
    opt = fused_optim.FusedSGD(model, lr=0.05, momentum=0.9, weight_decay=1e-4)
    for step in range(3):
        xb, yb = torch.randn_like(x), torch.randint(0, 10, y.shape)
        opt_ref.zero_grad()
        torch.nn.functional.cross_entropy(ref(xb), yb).backward()
        opt_ref.step()
        if step == 1:                                   # another component clears gradients its own way
            for p in model.parameters():
                p.grad = None
            assert not nv.grad_views_intact()
        opt.zero_grad()
        assert nv.grad_views_intact()
        loss = torch.nn.functional.cross_entropy(nv.autograd_forward(xb, True), yb)
        loss.backward()
        opt.step()
    ref_params = dict(ref.named_parameters())
    for name, p in model.named_parameters():
        a, b = p.data.flatten(), ref_params[name].data.flatten()
        rel = ((a - b).norm() / (b.norm() + 1e-12)).item()
        assert rel < 2e-2, f"{name}: relative L2 {rel:.3e}"
