"""Checkpoint save / resume in the torch ``state_dict`` format of the reference model.

The reference never saves anything (no ``torch.save`` anywhere); the contract is
the *format* its ``model.py`` implies: 58 entries, un-prefixed keys, fp32, conv weights OIHW
(SURVEY.md 2.9).  Internally the GPU path keeps KRSC weights in flat buffers; ``portable_state_dict``
converts at this boundary, and ``load`` accepts both plain and ``module.``-prefixed (DDP) keys.
"""
from __future__ import annotations

import os
from typing import Optional

import torch


def save_checkpoint(path: str, model, optimizer_state: Optional[dict] = None, step: int = 0) -> None:
    sd = model.portable_state_dict() if hasattr(model, "portable_state_dict") else \
        {k: v.detach().cpu().contiguous() for k, v in model.state_dict().items()}
    payload = {"model": sd, "optimizer": optimizer_state, "step": int(step), "format": "b200dp-v1"}
    tmp = path + ".tmp"
    torch.save(payload, tmp)
    os.replace(tmp, path)


def load_checkpoint(path: str, model, strict: bool = True) -> dict:
    # tensors, numbers and plain containers only: the restricted unpickler is enough and refuses code
    ck = torch.load(path, map_location="cpu", weights_only=True)
    sd = ck["model"] if isinstance(ck, dict) and "model" in ck else ck  # also accept a bare state_dict
    if hasattr(model, "load_portable_state_dict"):
        model.load_portable_state_dict(sd, strict=strict)
    else:
        sd = {(k[len("module."):] if k.startswith("module.") else k): v for k, v in sd.items()}
        model.load_state_dict(sd, strict=strict)
    return ck if isinstance(ck, dict) and "model" in ck else {"model": sd, "optimizer": None, "step": 0}
