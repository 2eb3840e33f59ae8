"""Native batch loader (csrc/loader.cpp + data.NativeBatchPipeline): host-only C++ threads, so everything
about it is checked on CPU -- transform parity with the torch implementation of the reference's
torchvision pipeline (ref: master/part2a/part2a.py:91-118), ordering / sharding, determinism, restarts."""
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.dirname(__file__))
from cs744_pytorch_distributed_tutorial_b200.data import (BatchPipeline, NativeBatchPipeline,  # noqa: E402
                                                          ShardedSampler, SyntheticCIFAR10, augment_batch)


@pytest.fixture(scope="module")
def ds():
    return SyntheticCIFAR10(1000, True)


def _normalised(images_u8: np.ndarray) -> torch.Tensor:
    out = torch.empty(len(images_u8), 3, 32, 32)
    augment_batch(torch.from_numpy(images_u8), out, None, False)
    return out


def test_eval_batches_are_bit_identical_to_the_torch_pipeline(ds):
    nat = NativeBatchPipeline(ds, 128, train=False, pin=False, threads=3)
    ref = BatchPipeline(ds, 128, train=False, pin=False)
    n = 0
    for (x, y), (xr, yr) in zip(nat, ref):
        assert torch.equal(x, xr) and torch.equal(y, yr)
        n += len(y)
    assert n == len(ds) and len(nat) == len(ref) == 8          # 7 full batches + the partial one
    nat.close()


def test_training_order_is_the_sharded_sampler_and_shards_are_disjoint(ds):
    seen = []
    for rank in range(2):
        p = NativeBatchPipeline(ds, 64, world=2, rank=rank, train=True, augment=False, pin=False)
        idx = ShardedSampler(len(ds), 2, rank, shuffle=True, seed=0).indices()
        ys = torch.cat([y.clone() for _, y in p])
        xs0 = next(iter(p))[0].clone()
        assert torch.equal(ys, torch.from_numpy(ds.targets)[idx])
        assert torch.equal(xs0, _normalised(ds.data[idx[:64].numpy()]))
        seen.append(set(idx.tolist()))
        p.close()
    assert seen[0].isdisjoint(seen[1]) and len(seen[0] | seen[1]) == len(ds)


def test_augmented_samples_are_crops_and_flips_of_their_source(ds):
    p = NativeBatchPipeline(ds, 50, train=True, augment=True, pin=False, threads=2)
    idx = p.sampler.indices()
    x, _ = next(iter(p))
    x = x.clone()
    found = set()
    for i in range(50):
        src = torch.from_numpy(ds.data[idx[i]]).permute(2, 0, 1)              # CHW uint8
        padded = torch.zeros(3, 40, 40, dtype=torch.uint8)
        padded[:, 4:36, 4:36] = src
        match = None
        for di in range(9):
            for dj in range(9):
                win = padded[:, di:di + 32, dj:dj + 32]
                for flip in (0, 1):
                    cand = win.flip(2) if flip else win
                    if torch.equal(_normalised(cand.permute(1, 2, 0).unsqueeze(0).numpy())[0], x[i]):
                        match = (di, dj, flip)
        assert match is not None, f"sample {i} is not a crop/flip of its source image"
        found.add(match)
    assert len(found) > 25                        # 162 possible draws: 50 samples must not collapse onto a few
    assert any(f for _, _, f in found) and any(not f for _, _, f in found)
    p.close()


def test_batches_do_not_depend_on_the_thread_count_and_change_with_the_epoch(ds):
    def epoch(threads, ep):
        p = NativeBatchPipeline(ds, 100, train=True, augment=True, pin=False, threads=threads, ring=6)
        p.sampler.set_epoch(ep)
        out = torch.cat([x.clone() for x, _ in p])
        p.close()
        return out
    a, b, c = epoch(1, 0), epoch(5, 0), epoch(5, 1)
    assert torch.equal(a, b)
    assert not torch.equal(b, c)


def test_abandoned_epoch_restart_and_drop_last(ds):
    p = NativeBatchPipeline(ds, 64, train=True, augment=True, pin=False, threads=4, drop_last=True)
    assert len(p) == 1000 // 64
    it = iter(p)
    first = [next(it)[1].clone() for _ in range(3)]          # stop early, slots still held
    again = [y.clone() for _, y in p]                        # a fresh epoch over the same order
    assert len(again) == len(p) and all(len(y) == 64 for y in again)
    assert all(torch.equal(a, b) for a, b in zip(first, again))
    p.close()
    p.close()                                                # idempotent


def test_loader_rejects_bad_arguments(ds):
    with pytest.raises(ValueError):
        NativeBatchPipeline(ds, 64, ring=3, release_lag=2, pin=False)
    bad = SyntheticCIFAR10(10, True)
    bad.data = bad.data[:, :16]
    with pytest.raises(ValueError):
        NativeBatchPipeline(bad, 4, pin=False)


def test_native_loader_is_much_faster_than_the_python_pipeline():
    import time
    big = SyntheticCIFAR10(8192, True)

    def rate(p):
        it = iter(p)
        next(it)
        t0, n = time.perf_counter(), 0
        for x, _ in it:
            n += len(x)
        return n / (time.perf_counter() - t0)
    slow = rate(BatchPipeline(big, 256, pin=False))
    nat = NativeBatchPipeline(big, 256, pin=False, threads=2)
    fast = rate(nat)
    nat.close()
    assert fast > 3 * slow, (fast, slow)


def test_thread_count_zero_means_auto_and_gives_the_same_batches(ds, monkeypatch):
    """``threads=0`` (the default of ``TrainConfig.loader_threads``): 2 x host CPUs / ranks, clamped to 4..8 --
    and, like any thread count, the same batches."""
    import os
    monkeypatch.setattr(os, "cpu_count", lambda: 32)
    auto8 = NativeBatchPipeline(ds, 100, world=8, rank=3, train=True, augment=True, pin=False, threads=0)
    assert auto8.threads == 8                                  # 2 * 32 / 8
    monkeypatch.setattr(os, "cpu_count", lambda: 4)
    auto4 = NativeBatchPipeline(ds, 100, world=8, rank=3, train=True, augment=True, pin=False, threads=0)
    assert auto4.threads == 4                                  # never below 4
    fixed = NativeBatchPipeline(ds, 100, world=8, rank=3, train=True, augment=True, pin=False, threads=2)
    for (xa, ya), (xb, yb), (xc, yc) in zip(auto8, auto4, fixed):
        assert torch.equal(xa, xc) and torch.equal(ya, yc) and torch.equal(xb, xc) and torch.equal(yb, yc)
    for p in (auto8, auto4, fixed):
        p.close()
