"""Runs the UNMODIFIED reference scripts from ``baseline/_ref`` for the benchmark's reference arm.

What is the reference's and what is harness:
  * reference code, executed as is: ``model.py`` (``VGG11``) and the ``train_model`` function of
    ``master/<part>/<script>.py`` (rank 0) or ``slave/...`` (other ranks) -- its loop, its
    ``.to(device)``, ``zero_grad``, forward, loss, backward, gradient-sync block, ``optimizer.step``
    and ``.item()`` prints.
  * harness, mirroring the few lines of the reference's ``vgg_model`` / ``init_process`` that cannot
    run here: process-group init (nccl instead of gloo), ``VGG11().to(device)``, ``DDP(model)`` for
    part3, ``optim.SGD(lr 0.1, momentum 0.9, wd 1e-4)``, ``CrossEntropyLoss`` -- and a loader that
    serves synthetic CIFAR-shaped batches and brackets exactly K iterations with CUDA events.
  * parameterisation applied IN MEMORY at load time (files on disk stay byte-identical, checked
    against the install manifest): the module global ``device`` -> "cuda"; for part2a / part2a_extra
    / part2b the hard-coded world size 4 (``[0, 1, 2, 3]``, ``range(4)``, ``/ 4``) -> the actual world size;
    for part2a_extra at a world size other than 4 the master's hand-unrolled peers 1, 2, 3 become a loop over
    ``range(1, world)`` with the same statements in the same order.  At world size 4
    every script runs byte-identical.  part3 needs no source parameterisation.
"""
from __future__ import annotations

import contextlib
import hashlib
import io
import json
import re
import sys
import types
from pathlib import Path
from typing import List, Optional, Tuple

import torch

REF_ROOT = Path(__file__).resolve().parent / "_ref"

SCRIPTS = {
    "part1": ("part1", "part1.py"),
    "part2a": ("part2a", "part2a.py"),
    "part2a_extra": ("part2a", "part2a_extra.py"),
    "part2b": ("part2b", "part2b.py"),
    "part3": ("part3", "part3.py"),
}


def available() -> Optional[str]:
    """None if the reference can run, else a one-line reason."""
    if not (REF_ROOT / "MANIFEST.json").exists():
        return "baseline/_ref is missing (run baseline/install_reference.py where /root/reference is mounted)"
    try:
        import torchvision  # noqa: F401  (the reference imports it at module level)
    except Exception as e:  # pragma: no cover
        return f"torchvision not importable: {e}"
    return None


def _verify(path: Path) -> None:
    manifest = json.loads((REF_ROOT / "MANIFEST.json").read_text())["sha256"]
    rel = str(path.relative_to(REF_ROOT))
    got = hashlib.sha256(path.read_bytes()).hexdigest()
    if manifest.get(rel) != got:
        raise RuntimeError(f"{rel} differs from the installed reference (sha256 mismatch)")


def _parameterise_world(src: str, world: int, script: str) -> str:
    """The reference hard-codes 4 ranks in its sync blocks; substitute the real world size."""
    if world == 4:
        return src
    src = src.replace("dist.new_group([0, 1, 2, 3])", f"dist.new_group(list(range({world})))")
    src = src.replace("range(4)", f"range({world})")
    src = re.sub(r"/ 4\b", f"/ {world}", src)
    if script == "part2a_extra.py" and "grad_1 = torch.zeros_like(p.grad)" in src:
        # the master unrolls peers 1, 2, 3 by hand: the same statements as a loop over range(1, world), in the
        # same order (blocking irecv per peer, left-to-right sum starting from the local gradient, one true
        # division, blocking isend per peer) -- SURVEY.md 6.3-4.  The slave script has no world-size literal.
        a = src.index("            grad_1 = torch.zeros_like(p.grad)")
        b = src.index("        optimizer.step()", a)
        body = (
            f"            grads = [torch.zeros_like(p.grad) for _ in range({world} - 1)]\n"
            f"            for peer in range(1, {world}):\n"
            "                req = dist.irecv(tensor=grads[peer - 1], src=peer)\n"
            "                req.wait()\n"
            "            grad_mean = p.grad\n"
            "            for g in grads:\n"
            "                grad_mean = grad_mean + g\n"
            f"            grad_mean = grad_mean / {world}\n"
            "            p.grad = grad_mean\n"
            f"            for peer in range(1, {world}):\n"
            "                req = dist.isend(tensor=grad_mean, dst=peer)\n"
            "                req.wait()\n")
        src = src[:a] + body + src[b:]
    return src


def load_reference(mode: str, rank: int, world: int, device: str) -> types.ModuleType:
    part, script = SCRIPTS[mode]
    role = "master" if (rank == 0 or mode == "part1") else "slave"
    path = REF_ROOT / role / part / script
    _verify(path)
    _verify(path.parent / "model.py")
    src = _parameterise_world(path.read_text(), world, script)
    sys.dont_write_bytecode = True
    sys.path.insert(0, str(path.parent))
    sys.modules.pop("model", None)
    try:
        mod = types.ModuleType(f"reference_{mode}_{role}")
        mod.__file__ = str(path)
        exec(compile(src, str(path), "exec"), mod.__dict__)
    finally:
        sys.path.remove(str(path.parent))
    mod.device = device
    return mod


class TimedLoader:
    """Serves ``warmup + steps`` batches and brackets the last ``steps`` with CUDA events.

    ``__next__`` of iteration i runs after iteration i-1 has been fully enqueued, so recording the
    start event when batch ``warmup`` is requested and the end event when the loader is exhausted
    times exactly ``steps`` iterations of the reference's own loop.
    """

    def __init__(self, batches: List[Tuple[torch.Tensor, torch.Tensor]], warmup: int, steps: int, world: int):
        self.batches, self.warmup, self.steps, self.world = batches, warmup, steps, world
        self.start = torch.cuda.Event(enable_timing=True)
        self.end = torch.cuda.Event(enable_timing=True)
        self.elapsed_ms: Optional[float] = None
        self.on_timed_region = None  # optional callbacks (clock sampling)
        self.on_timed_region_end = None

    def __len__(self) -> int:
        return self.warmup + self.steps

    def __iter__(self):
        import torch.distributed as dist
        source = None if isinstance(self.batches, list) else iter(self.batches)     # a DataLoader
        for i in range(self.warmup + self.steps):
            if i == self.warmup:
                if self.world > 1:
                    dist.barrier()
                torch.cuda.synchronize()
                if self.on_timed_region:
                    self.on_timed_region()
                self.start.record()
            yield self.batches[i % len(self.batches)] if source is None else next(source)
        del source
        self.end.record()
        torch.cuda.synchronize()
        if self.on_timed_region_end:
            self.on_timed_region_end()
        self.elapsed_ms = self.start.elapsed_time(self.end)


def reference_dataloader(per_rank_batch: int, rank: int, world: int, n_images: int):
    """The reference's input pipeline on synthetic pixels: torchvision's ``CIFAR10`` dataset class (its own
    ``__getitem__``: uint8 array -> PIL image -> transform) with the tutorial's transform chain and its
    ``DataLoader(num_workers=2, pin_memory=True)`` (ref: master/part1/part1.py:65-85; DistributedSampler for
    world > 1, ref: master/part3/part3.py:94-102).  Only the pixel source differs: the CIFAR-10 files are not on
    the box and cannot be downloaded, so the arrays the class would have unpickled are random."""
    import numpy as np
    from torch.utils.data import DataLoader
    from torch.utils.data.distributed import DistributedSampler
    from torchvision import datasets, transforms
    rng = np.random.default_rng(1234)
    ds = datasets.CIFAR10.__new__(datasets.CIFAR10)       # skips the download / md5 check of __init__ only
    ds.root, ds.train = "", True
    ds.data = rng.integers(0, 256, size=(n_images, 32, 32, 3), dtype=np.uint8)
    ds.targets = rng.integers(0, 10, size=n_images).tolist()
    ds.transform = transforms.Compose([
        transforms.RandomCrop(32, padding=4), transforms.RandomHorizontalFlip(), transforms.ToTensor(),
        transforms.Normalize(mean=[v / 255.0 for v in (125.3, 123.0, 113.9)],
                             std=[v / 255.0 for v in (63.0, 62.1, 66.7)])])
    ds.target_transform, ds.transforms = None, None
    if world > 1:
        sampler = DistributedSampler(ds, num_replicas=world, rank=rank)
        return DataLoader(ds, num_workers=2, batch_size=per_rank_batch, sampler=sampler, pin_memory=True)
    return DataLoader(ds, num_workers=2, batch_size=per_rank_batch, sampler=None, shuffle=True, pin_memory=True)


def run_reference(mode: str, rank: int, world: int, local_rank: int, per_rank_batch: int, steps: int, warmup: int,
                  host_inputs: bool, n_distinct_batches: int = 8, quiet: bool = True,
                  on_timed_region=None, on_timed_region_end=None, real_loader: bool = False) -> dict:
    """Time ``steps`` iterations of the reference's ``train_model``.  Returns ms per step (this rank)."""
    import torch.distributed as dist
    import torch.nn as nn
    import torch.optim as optim
    torch.cuda.set_device(local_rank)
    device = "cuda"
    if world > 1 and not dist.is_initialized():
        dist.init_process_group("nccl", rank=rank, world_size=world)
    ref = load_reference(mode, rank, world, device)
    torch.manual_seed(5000)            # ref: master/part3/part3.py:78
    model = ref.mdl.VGG11()
    model.to(device)
    if mode == "part3":
        model = ref.DDP(model)         # ref: master/part3/part3.py:116
    optimizer = optim.SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=0.0001)
    criterion = nn.CrossEntropyLoss().to(device)
    g = torch.Generator().manual_seed(1234 + rank)
    batches = []
    if real_loader:
        batches = reference_dataloader(per_rank_batch, rank, world, (warmup + steps) * per_rank_batch * world)
    for _ in range(0 if real_loader else n_distinct_batches):
        x = torch.randn(per_rank_batch, 3, 32, 32, generator=g)
        y = torch.randint(0, 10, (per_rank_batch,), generator=g)
        if host_inputs:
            batches.append((x.pin_memory(), y.pin_memory()))
        else:
            batches.append((x.cuda(), y.cuda()))
    loader = TimedLoader(batches, warmup, steps, world)
    loader.on_timed_region, loader.on_timed_region_end = on_timed_region, on_timed_region_end
    sink = io.StringIO()
    with contextlib.redirect_stdout(sink) if quiet else contextlib.nullcontext():
        if mode == "part1":
            ref.train_model(model, loader, optimizer, criterion, 0)     # ref: master/part1/part1.py:20
        else:
            ref.train_model(model, loader, optimizer, criterion, 0, rank)
    ms = loader.elapsed_ms / steps
    h2d = per_rank_batch * (3 * 32 * 32 * 4 + 8) if host_inputs else 0
    return {"ms_per_step": ms, "h2d_bytes_per_step": h2d, "stdout_tail": sink.getvalue()[-400:]}
