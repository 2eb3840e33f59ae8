"""VGG for 3x32x32 inputs / 10 classes with the reference's module tree and state_dict.

The module tree is ``layers`` (a ``Sequential`` of Conv2d / BatchNorm2d / ReLU /
MaxPool2d) plus ``fc1``, so parameter order, names (``layers.0.weight`` ...
``fc1.bias``) and the 58-entry state_dict are those of the reference
(ref: master/part1/model.py:1-50).  On CPU the forward pass runs the
stock torch ops (the gloo plumbing path and the numerics oracle); on a CUDA device
``enable_native()`` swaps the whole forward/backward for the sm_100a executor in
``native_vgg.py`` while the module keeps owning the parameters.
"""
from __future__ import annotations

from typing import List

import torch
import torch.nn as nn

# channel plan per architecture; 'M' is a 2x2/2 max-pool (ref: master/part1/model.py:3-8)
ARCH_PLANS = {
    "VGG11": (64, "M", 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"),
    "VGG13": (64, 64, "M", 128, 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"),
    "VGG16": (64, 64, "M", 128, 128, "M", 256, 256, 256, "M", 512, 512, 512, "M", 512, 512, 512, "M"),
    "VGG19": (64, 64, "M", 128, 128, "M", 256, 256, 256, 256, "M", 512, 512, 512, 512, "M",
              512, 512, 512, 512, "M"),
}
NUM_CLASSES = 10
FLATTEN_FEATURES = 512


def build_feature_stack(plan) -> nn.Sequential:
    mods: List[nn.Module] = []
    cin = 3
    for item in plan:
        if item == "M":
            mods.append(nn.MaxPool2d(kernel_size=2, stride=2))
            continue
        mods += [nn.Conv2d(cin, item, kernel_size=3, stride=1, padding=1, bias=True),
                 nn.BatchNorm2d(item), nn.ReLU(inplace=True)]
        cin = item
    return nn.Sequential(*mods)


class ConvBlockSpec:
    """One conv -> BN -> ReLU (-> pool) block, as indices into ``VGG.layers``."""

    def __init__(self, conv_idx: int, cin: int, cout: int, hw: int, pool: bool):
        self.conv_idx, self.bn_idx = conv_idx, conv_idx + 1
        self.cin, self.cout, self.hw, self.pool = cin, cout, hw, pool

    def __repr__(self) -> str:
        return f"ConvBlockSpec(layers.{self.conv_idx}: {self.cin}->{self.cout} @{self.hw}x{self.hw} pool={self.pool})"


def block_specs(plan) -> List[ConvBlockSpec]:
    specs: List[ConvBlockSpec] = []
    idx, cin, hw = 0, 3, 32
    for item in plan:
        if item == "M":
            specs[-1].pool = True
            hw //= 2
            idx += 1
        else:
            specs.append(ConvBlockSpec(idx, cin, item, hw, False))
            cin = item
            idx += 3
    return specs


class VGG(nn.Module):
    def __init__(self, arch: str = "VGG11"):
        super().__init__()
        if arch not in ARCH_PLANS:
            raise ValueError(f"unknown architecture {arch!r}; choose from {sorted(ARCH_PLANS)}")
        self.arch = arch
        self.layers = build_feature_stack(ARCH_PLANS[arch])
        self.fc1 = nn.Linear(FLATTEN_FEATURES, NUM_CLASSES)
        self._native = None  # NativeVGG executor once enable_native() was called

    # -- structure helpers ---------------------------------------------------------
    def block_specs(self) -> List[ConvBlockSpec]:
        return block_specs(ARCH_PLANS[self.arch])

    # -- native path -----------------------------------------------------------------
    def enable_native(self, max_batch: int, heap=None, **kw):
        """Move parameters into flat (optionally peer-visible) device buffers and route
        forward/backward through the sm_100a kernels.  Requires the module on a CUDA device."""
        from .native_vgg import NativeVGG
        self._native = NativeVGG(self, max_batch=max_batch, heap=heap, **kw)
        return self._native

    @property
    def native(self):
        return self._native

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self._native is not None and x.is_cuda:
            return self._native.autograd_forward(x, self.training)
        y = self.layers(x)
        y = y.view(y.size(0), -1)
        return self.fc1(y)

    # -- checkpoint helpers: always un-prefixed, fp32, OIHW ----------------------------
    def portable_state_dict(self) -> dict:
        return {k: v.detach().to("cpu").contiguous().clone() for k, v in self.state_dict().items()}

    def load_portable_state_dict(self, sd: dict, strict: bool = True):
        sd = {(k[len("module."):] if k.startswith("module.") else k): v for k, v in sd.items()}
        with torch.no_grad():
            missing = self.load_state_dict(sd, strict=strict)
        if self._native is not None:
            self._native.refresh_from_module()
        return missing


def VGG11() -> VGG:
    """Factory with the reference's name (ref: master/part1/model.py:49-50)."""
    return VGG("VGG11")


def VGG13() -> VGG:
    return VGG("VGG13")


def VGG16() -> VGG:
    return VGG("VGG16")


def VGG19() -> VGG:
    return VGG("VGG19")
