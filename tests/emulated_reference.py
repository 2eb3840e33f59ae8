"""Plain-PyTorch fp32 reference of the *same op sequence* as the native executor: identical math,
with bf16 rounding inserted exactly where the kernels store bf16 (conv operands, conv outputs,
block outputs and the corresponding gradients).  ReLU masks and pool arg-max decisions are
discontinuous, so a pure-fp32 model and a bf16 pipeline legitimately disagree on a few percent of
them; rounding at the same points removes that ambiguity and lets gradients be compared tightly.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F


def rbf(x: torch.Tensor) -> torch.Tensor:
    """Round to bf16 with a straight-through gradient."""
    return x + (x.bfloat16().float() - x).detach()


def _round_grad(t: torch.Tensor) -> torch.Tensor:
    if t.requires_grad:
        t.register_hook(lambda g: g.bfloat16().float())
    return t


def emulated_forward(model: nn.Module, x: torch.Tensor, training: bool = True) -> torch.Tensor:
    """``model`` is a stock-torch VGG (fp32 parameters); returns logits."""
    h = x
    mods = list(model.layers)
    i = 0
    while i < len(mods):
        conv, bn = mods[i], mods[i + 1]
        assert isinstance(conv, nn.Conv2d) and isinstance(bn, nn.BatchNorm2d)
        i += 3
        pool = i < len(mods) and isinstance(mods[i], nn.MaxPool2d)
        if pool:
            i += 1
        y = F.conv2d(rbf(h), rbf(conv.weight), conv.bias, padding=1)
        y = _round_grad(rbf(y))
        z = F.batch_norm(y, bn.running_mean, bn.running_var, bn.weight, bn.bias, training=training,
                         momentum=bn.momentum, eps=bn.eps)
        if training:
            bn.num_batches_tracked += 1
        a = F.relu(z)
        o = F.max_pool2d(a, 2, 2) if pool else a
        h = _round_grad(rbf(o))
    return F.linear(h.flatten(1), model.fc1.weight, model.fc1.bias)
