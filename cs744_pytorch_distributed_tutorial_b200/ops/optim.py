"""``torch.optim``-style face of the fused SGD kernel for the eager / autograd API.

The reference builds ``optim.SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=0.0001)`` and
calls ``zero_grad()`` / ``step()`` around ``loss.backward()`` (ref: master/part1/part1.py:34,38,98-99).
With a native-enabled model the same three calls work, but ``step()`` is ONE kernel over the flat
parameter / gradient / momentum buffers followed by the refresh of the bf16 operand copies, and
``zero_grad()`` keeps ``p.grad`` pointing into the flat gradient buffer (never sets it to None).
"""
from __future__ import annotations

import torch

from . import native as K


class FusedSGD(torch.optim.Optimizer):
    def __init__(self, model, lr: float = 0.1, momentum: float = 0.9, weight_decay: float = 1e-4):
        nv = getattr(model, "native", None)
        if nv is None:
            raise ValueError("FusedSGD needs a model on which enable_native() was called")
        self.native = nv
        super().__init__(list(model.parameters()), dict(lr=lr, momentum=momentum, weight_decay=weight_decay))

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        g = self.param_groups[0]
        nv = self.native
        if not nv.grad_views_intact():
            # someone replaced p.grad (e.g. zero_grad(set_to_none=True) of another optimizer): copy back in
            params = dict(nv.module.named_parameters())
            for e in nv.layout.entries:
                p = params[e.name]
                if p.grad is not None:
                    nv._view(nv.flat_grads, e).copy_(p.grad)
            nv.restore_grad_views()
        K.sgd_momentum(nv.flat_params, nv.flat_grads, nv.flat_momentum, g["lr"], g["momentum"], g["weight_decay"])
        nv.refresh_shadows()
        return loss

    def zero_grad(self, set_to_none: bool = False) -> None:
        self.native.flat_grads.zero_()
        if not self.native.grad_views_intact():
            self.native.restore_grad_views()
