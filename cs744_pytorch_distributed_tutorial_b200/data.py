"""Data pipeline: CIFAR-10-shaped datasets, the reference's augmentation, rank sharding.

The reference builds torchvision ``CIFAR10`` with ``RandomCrop(32, padding=4)``,
``RandomHorizontalFlip``, ``ToTensor`` and ``Normalize`` and shards it with
``DistributedSampler`` behind a 2-worker ``DataLoader``
(ref: master/part2a/part2a.py:91-122).  There is no network here, so the default
dataset is a deterministic synthetic one of the same shape (uint8 HWC images, 10
classes); real CIFAR-10 batches on disk are used when present.

Two consumers:
  * ``make_loaders`` -- ``torch.utils.data.DataLoader`` objects with the reference's
    settings, per-sample transforms; API parity with the tutorial.
  * ``BatchPipeline`` -- vectorised whole-batch augmentation into pinned, double
    buffered host tensors; this is what feeds the GPU engine so that the input side
    is not the bottleneck of a ~1 ms training step.
"""
from __future__ import annotations

import math
import os
import pickle
from typing import Iterator, Optional, Tuple

import numpy as np
import torch
from torch.utils.data import DataLoader, Dataset, Sampler

CIFAR_MEAN = tuple(x / 255.0 for x in (125.3, 123.0, 113.9))  # ref: master/part2a/part2a.py:91
CIFAR_STD = tuple(x / 255.0 for x in (63.0, 62.1, 66.7))      # ref: master/part2a/part2a.py:92


# ---------------------------------------------------------------------------
# datasets
# ---------------------------------------------------------------------------
class SyntheticCIFAR10(Dataset):
    """``n`` random uint8 32x32x3 images with random labels 0..9, fixed by ``seed``.

    Images carry a weak class-dependent signal (a per-class mean colour offset) so that
    a training run can actually reduce the loss, which makes convergence smoke tests
    meaningful without the real dataset.
    """

    def __init__(self, n: int, train: bool = True, seed: int = 1234, transform=None):
        rng = np.random.default_rng(seed + (0 if train else 1))
        self.targets = rng.integers(0, 10, size=n, dtype=np.int64)
        base = rng.integers(0, 256, size=(n, 32, 32, 3), dtype=np.int16)
        class_offset = (np.random.default_rng(99).integers(-40, 41, size=(10, 1, 1, 3))).astype(np.int16)
        self.data = np.clip(base + class_offset[self.targets], 0, 255).astype(np.uint8)
        self.train = train
        self.transform = transform

    def __len__(self) -> int:
        return len(self.targets)

    def __getitem__(self, i: int):
        img = torch.from_numpy(self.data[i])  # HWC uint8
        if self.transform is not None:
            img = self.transform(img)
        return img, int(self.targets[i])


class CIFAR10Batches(Dataset):
    """Real CIFAR-10 from the python-pickle batches, if ``root/cifar-10-batches-py`` exists."""

    def __init__(self, root: str, train: bool = True, transform=None):
        d = os.path.join(root, "cifar-10-batches-py")
        names = [f"data_batch_{i}" for i in range(1, 6)] if train else ["test_batch"]
        xs, ys = [], []
        for nme in names:
            with open(os.path.join(d, nme), "rb") as f:
                entry = pickle.load(f, encoding="latin1")
            xs.append(entry["data"])
            ys.extend(entry["labels"])
        self.data = np.vstack(xs).reshape(-1, 3, 32, 32).transpose(0, 2, 3, 1).copy()
        self.targets = np.asarray(ys, dtype=np.int64)
        self.train = train
        self.transform = transform

    @staticmethod
    def present(root: str) -> bool:
        return os.path.isdir(os.path.join(root, "cifar-10-batches-py"))

    def __len__(self) -> int:
        return len(self.targets)

    def __getitem__(self, i: int):
        img = torch.from_numpy(self.data[i])
        if self.transform is not None:
            img = self.transform(img)
        return img, int(self.targets[i])


def build_dataset(cfg, train: bool, transform=None) -> Dataset:
    if not cfg.synthetic:
        if not CIFAR10Batches.present(cfg.data_root):
            # a silent fall-back would report losses and accuracies that look like CIFAR-10 results and are not
            raise FileNotFoundError(
                f"real data was requested but {os.path.join(cfg.data_root, 'cifar-10-batches-py')} does not exist "
                "(there is no download path here); drop --real-data to train on the synthetic set")
        return CIFAR10Batches(cfg.data_root, train, transform)
    n = cfg.train_size if train else cfg.test_size
    return SyntheticCIFAR10(n, train, seed=1234, transform=transform)


# ---------------------------------------------------------------------------
# transforms (tensor implementations of the torchvision ops the reference composes)
# ---------------------------------------------------------------------------
class TrainTransform:
    """RandomCrop(32, padding=4) + RandomHorizontalFlip + ToTensor + Normalize on one HWC uint8 image."""

    def __init__(self, augment: bool = True):
        self.augment = augment
        self.mean = torch.tensor(CIFAR_MEAN).view(3, 1, 1)
        self.std = torch.tensor(CIFAR_STD).view(3, 1, 1)

    def __call__(self, img: torch.Tensor) -> torch.Tensor:
        x = img.permute(2, 0, 1)
        if self.augment:
            padded = torch.zeros(3, 40, 40, dtype=x.dtype)
            padded[:, 4:36, 4:36] = x
            i, j = torch.randint(0, 9, (2,)).tolist()
            x = padded[:, i:i + 32, j:j + 32]
            if torch.rand(()) < 0.5:
                x = x.flip(2)
        return (x.float().div_(255.0) - self.mean) / self.std


class TestTransform(TrainTransform):
    def __init__(self):
        super().__init__(augment=False)


def augment_batch(images_u8: torch.Tensor, out: torch.Tensor, gen: Optional[torch.Generator], augment: bool) -> None:
    """Vectorised version of the transform for a whole ``[B, 32, 32, 3]`` uint8 batch.

    Writes normalised fp32 NCHW into ``out`` (which may be pinned memory).
    """
    B = images_u8.shape[0]
    x = images_u8.permute(0, 3, 1, 2)
    if augment:
        padded = torch.zeros(B, 3, 40, 40, dtype=torch.uint8)
        padded[:, :, 4:36, 4:36] = x
        di = torch.randint(0, 9, (B,), generator=gen)
        dj = torch.randint(0, 9, (B,), generator=gen)
        flip = torch.rand(B, generator=gen) < 0.5
        rows = di.view(B, 1) + torch.arange(32).view(1, 32)                     # [B,32]
        cols = dj.view(B, 1) + torch.arange(32).view(1, 32)
        cols = torch.where(flip.view(B, 1), cols.flip(1), cols)
        bidx = torch.arange(B).view(B, 1, 1)
        x = padded.permute(0, 2, 3, 1)[bidx, rows.view(B, 32, 1), cols.view(B, 1, 32)]  # [B,32,32,3]
        x = x.permute(0, 3, 1, 2)
    mean = torch.tensor(CIFAR_MEAN).view(1, 3, 1, 1)
    std = torch.tensor(CIFAR_STD).view(1, 3, 1, 1)
    torch.div(x.float(), 255.0, out=out)
    out.sub_(mean).div_(std)


# ---------------------------------------------------------------------------
# sharding
# ---------------------------------------------------------------------------
class ShardedSampler(Sampler):
    """Rank shard of a dataset with ``DistributedSampler`` semantics.

    Shuffles with ``seed + epoch``, pads by wrapping around to a multiple of the world
    size, then takes every ``world``-th index starting at ``rank``
    (ref: master/part2a/part2a.py:107 uses ``DistributedSampler(training_set)``: shuffle,
    seed 0, no ``set_epoch`` because there is a single epoch).
    """

    def __init__(self, n: int, world: int, rank: int, shuffle: bool = True, seed: int = 0, drop_last: bool = False):
        self.n, self.world, self.rank, self.shuffle, self.seed, self.drop_last = n, world, rank, shuffle, seed, drop_last
        self.epoch = 0
        if drop_last and n % world:
            self.num_samples = n // world
        else:
            self.num_samples = math.ceil(n / world)
        self.total = self.num_samples * world

    def set_epoch(self, epoch: int) -> None:
        self.epoch = epoch

    def indices(self) -> torch.Tensor:
        if self.shuffle:
            g = torch.Generator()
            g.manual_seed(self.seed + self.epoch)
            idx = torch.randperm(self.n, generator=g)
        else:
            idx = torch.arange(self.n)
        if self.drop_last:
            idx = idx[: self.total]
        elif self.total > self.n:
            pad = self.total - self.n
            reps = math.ceil(pad / self.n)
            idx = torch.cat([idx, idx.repeat(reps)[:pad]])
        return idx[self.rank: self.total: self.world]

    def __iter__(self) -> Iterator[int]:
        return iter(self.indices().tolist())

    def __len__(self) -> int:
        return self.num_samples


def make_loaders(cfg) -> Tuple[DataLoader, DataLoader]:
    """DataLoaders with the reference's settings (ref: master/part2a/part2a.py:103-122): sharded,
    shuffled training set; full, unshuffled test set on every rank; 2 workers; pinned memory."""
    train_set = build_dataset(cfg, True, TrainTransform(cfg.augment))
    test_set = build_dataset(cfg, False, TestTransform())
    sampler = ShardedSampler(len(train_set), cfg.world_size, cfg.rank, shuffle=True, seed=0, drop_last=False)
    pin = torch.cuda.is_available()
    train_loader = DataLoader(train_set, batch_size=cfg.per_rank_batch, sampler=sampler, num_workers=cfg.num_workers,
                              pin_memory=pin, drop_last=cfg.drop_last)
    test_loader = DataLoader(test_set, batch_size=cfg.per_rank_batch, shuffle=False, num_workers=cfg.num_workers,
                             pin_memory=pin)
    return train_loader, test_loader


class BatchPipeline:
    """Whole-batch loader: gathers a shard's uint8 images, augments them with vectorised
    tensor ops and hands out pinned fp32 NCHW batches from a small ring of buffers."""

    def __init__(self, dataset, batch: int, world: int = 1, rank: int = 0, train: bool = True, augment: bool = True,
                 drop_last: bool = False, seed: int = 0, ring: int = 3, pin: Optional[bool] = None):
        self.data = torch.from_numpy(dataset.data)
        self.targets = torch.from_numpy(dataset.targets)
        self.batch, self.train, self.augment, self.drop_last = batch, train, augment and train, drop_last
        self.sampler = ShardedSampler(len(dataset), world if train else 1, rank if train else 0, shuffle=train,
                                      seed=seed, drop_last=False)
        pin = torch.cuda.is_available() if pin is None else pin
        self._x = [torch.empty(batch, 3, 32, 32, pin_memory=pin) for _ in range(ring)]
        self._y = [torch.empty(batch, dtype=torch.int64, pin_memory=pin) for _ in range(ring)]
        self._gen = torch.Generator().manual_seed(seed * 7919 + rank)
        self._slot = 0
        self._copied = [None] * ring      # per slot: event behind the host->device copy of its last batch
        self._last_slot = None

    def __len__(self) -> int:
        n = len(self.sampler)
        return n // self.batch if self.drop_last else math.ceil(n / self.batch)

    @property
    def dataset_len(self) -> int:
        return self.data.shape[0]

    def consumed(self, event=None) -> None:
        """The batch yielded last has been handed to the device; ``event`` (recorded behind its asynchronous
        host->device copy, or None) must have completed before that pinned buffer is written again.  Without
        this hand-shake a GPU running ``ring`` steps behind the host would copy a half-overwritten batch."""
        if self._last_slot is not None:
            self._copied[self._last_slot] = event

    def __iter__(self):
        idx = self.sampler.indices()
        for b in range(len(self)):
            sel = idx[b * self.batch: (b + 1) * self.batch]
            n = sel.numel()
            ev = self._copied[self._slot]
            if ev is not None:
                ev.synchronize()
                self._copied[self._slot] = None
            x, y = self._x[self._slot][:n], self._y[self._slot][:n]
            self._last_slot = self._slot
            self._slot = (self._slot + 1) % len(self._x)
            augment_batch(self.data[sel], x, self._gen, self.augment)
            torch.index_select(self.targets, 0, sel, out=y)
            yield x, y


class NativeBatchPipeline:
    """``BatchPipeline`` with the batch assembly in C++ threads (csrc/loader.cpp): crop, flip, uint8 -> fp32
    and normalisation are one table-lookup pass per image, ``ring`` batches are built ahead in order into
    pinned buffers.  Same sampling as ``BatchPipeline`` (``ShardedSampler``); the augmentation draws come
    from a counter-based generator keyed by (seed, epoch, position), so a run is reproducible for any
    thread count (they are not the draws of the torch generator ``BatchPipeline`` uses).

    A yielded batch stays valid until ``release_lag`` further batches have been requested; with a GPU
    consumer call ``consumed(event)`` after issuing the host->device copy of the batch, and the slot is
    only refilled once that event has completed."""

    def __init__(self, dataset, batch: int, world: int = 1, rank: int = 0, train: bool = True, augment: bool = True,
                 drop_last: bool = False, seed: int = 0, ring: int = 4, threads: int = 4, pin: Optional[bool] = None,
                 release_lag: int = 2):
        import collections
        import ctypes
        from .ops import _lib
        if ring < release_lag + 2:
            raise ValueError("ring must be at least release_lag + 2")
        if threads <= 0:                      # auto: oversubscribing the cores 2x measured best on an 8-GPU box
            import os
            threads = max(4, min(8, 2 * (os.cpu_count() or 8) // max(world, 1)))
        self.threads = threads
        self._lib, self._ct = _lib.lib(), ctypes
        self._images = np.ascontiguousarray(dataset.data, dtype=np.uint8)
        self._labels = np.ascontiguousarray(dataset.targets, dtype=np.int64)
        if self._images.shape[1:] != (32, 32, 3):
            raise ValueError(f"expected uint8 images [N,32,32,3], got {self._images.shape}")
        self.batch, self.train, self.augment, self.drop_last = batch, train, augment and train, drop_last
        self.sampler = ShardedSampler(len(dataset), world if train else 1, rank if train else 0, shuffle=train,
                                      seed=seed, drop_last=False)
        pin = torch.cuda.is_available() if pin is None else pin
        self._x = [torch.empty(batch, 3, 32, 32, pin_memory=pin) for _ in range(ring)]
        self._y = [torch.empty(batch, dtype=torch.int64, pin_memory=pin) for _ in range(ring)]
        vp = ctypes.c_void_p
        xs = (vp * ring)(*[t.data_ptr() for t in self._x])
        ys = (vp * ring)(*[t.data_ptr() for t in self._y])
        mean = (ctypes.c_float * 3)(*CIFAR_MEAN)
        std = (ctypes.c_float * 3)(*CIFAR_STD)
        h = vp()
        L = self._lib
        L.b200dp_loader_create.argtypes = [vp, vp, ctypes.c_longlong, ctypes.c_int, ctypes.c_int, vp, vp, vp, vp,
                                           ctypes.c_int, ctypes.c_ulonglong, ctypes.c_int, ctypes.POINTER(vp)]
        L.b200dp_loader_start_epoch.argtypes = [vp, vp, ctypes.c_longlong, ctypes.c_longlong, ctypes.c_int]
        L.b200dp_loader_next.argtypes = [vp, ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_int)]
        L.b200dp_loader_release.argtypes = [vp, ctypes.c_int]
        L.b200dp_loader_destroy.argtypes = [vp]
        self._check(L.b200dp_loader_create(self._images.ctypes.data, self._labels.ctypes.data, len(self._labels), batch,
                                           ring, xs, ys, mean, std, int(self.augment), seed * 7919 + rank, threads,
                                           ctypes.byref(h)), "b200dp_loader_create")
        self._h = h
        self.release_lag = release_lag
        self._held = collections.deque()
        self._pending_event = None

    def _check(self, rc: int, what: str) -> None:
        if rc != 0:
            from .ops._lib import NativeError
            msg = self._lib.b200dp_last_error()
            raise NativeError(f"{what} failed (code {rc}): {msg.decode() if msg else ''}")

    def __len__(self) -> int:
        n = len(self.sampler)
        return n // self.batch if self.drop_last else math.ceil(n / self.batch)

    @property
    def dataset_len(self) -> int:
        return self._images.shape[0]

    def consumed(self, event=None) -> None:
        """The batch yielded last has been handed to the device; ``event`` (a CUDA event recorded after its
        host->device copy, or None) must have completed before its buffer is refilled."""
        self._pending_event = event

    def _release_oldest(self) -> None:
        slot, ev = self._held.popleft()
        if ev is not None:
            ev.synchronize()
        self._check(self._lib.b200dp_loader_release(self._h, slot), "b200dp_loader_release")

    def __iter__(self):
        ct = self._ct
        order = np.ascontiguousarray(self.sampler.indices().numpy(), dtype=np.int64)
        while self._held:                      # an abandoned epoch: wait for its copies, hand the slots back
            self._release_oldest()
        self._check(self._lib.b200dp_loader_start_epoch(self._h, order.ctypes.data, len(order), self.sampler.epoch,
                                                        int(self.drop_last)), "b200dp_loader_start_epoch")
        slot, count = ct.c_int(), ct.c_int()
        while True:
            rc = self._lib.b200dp_loader_next(self._h, ct.byref(slot), ct.byref(count))
            if rc == 1:
                break
            self._check(rc, "b200dp_loader_next")
            self._pending_event = None
            yield self._x[slot.value][: count.value], self._y[slot.value][: count.value]
            self._held.append((slot.value, self._pending_event))
            while len(self._held) > self.release_lag:
                self._release_oldest()

    def close(self) -> None:
        if getattr(self, "_h", None) is not None:
            self._lib.b200dp_loader_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
