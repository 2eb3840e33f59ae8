"""CPU tests: model / state_dict parity with the reference, data pipeline, flat layout, config."""
import math
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.dirname(__file__))
from helpers import reference_model_module  # noqa: E402

from cs744_pytorch_distributed_tutorial_b200.config import TrainConfig  # noqa: E402
from cs744_pytorch_distributed_tutorial_b200.data import (BatchPipeline, ShardedSampler, SyntheticCIFAR10,  # noqa: E402
                                                          TrainTransform, augment_batch, make_loaders)
from cs744_pytorch_distributed_tutorial_b200.models.native_vgg import FlatLayout  # noqa: E402
from cs744_pytorch_distributed_tutorial_b200.models.vgg import VGG, VGG11, block_specs, ARCH_PLANS  # noqa: E402


def test_state_dict_contract_58_entries():
    sd = VGG11().state_dict()
    assert len(sd) == 58
    assert list(sd)[:7] == ["layers.0.weight", "layers.0.bias", "layers.1.weight", "layers.1.bias",
                            "layers.1.running_mean", "layers.1.running_var", "layers.1.num_batches_tracked"]
    assert sd["layers.25.weight"].shape == (512, 512, 3, 3) and sd["fc1.weight"].shape == (10, 512)
    assert sd["layers.1.num_batches_tracked"].dtype == torch.int64
    params = list(VGG11().named_parameters())
    assert len(params) == 34 and sum(p.numel() for _, p in params) == 9_231_114


def test_same_seed_gives_reference_weights_and_outputs():
    ref_mod = reference_model_module()
    if ref_mod is None:
        pytest.skip("reference model.py not available")
    torch.manual_seed(5000)
    ref = ref_mod.VGG11()
    torch.manual_seed(5000)
    mine = VGG11()
    rs, ms = ref.state_dict(), mine.state_dict()
    assert list(rs) == list(ms)
    for k in rs:
        assert torch.equal(rs[k], ms[k]), k
    x = torch.randn(4, 3, 32, 32)
    ref.eval(); mine.eval()
    assert torch.equal(ref(x), mine(x))
    # checkpoints move both ways with strict=True
    ref.load_state_dict(mine.portable_state_dict(), strict=True)
    mine.load_portable_state_dict({"module." + k: v for k, v in rs.items()}, strict=True)


@pytest.mark.parametrize("arch", sorted(ARCH_PLANS))
def test_all_archs_forward(arch):
    m = VGG(arch).eval()
    assert m(torch.randn(2, 3, 32, 32)).shape == (2, 10)
    specs = block_specs(ARCH_PLANS[arch])
    assert specs[0].cin == 3 and specs[-1].hw == 2 and specs[-1].pool
    assert sum(s.pool for s in specs) == 5


@pytest.mark.parametrize("n,world", [(50000, 4), (1000, 3), (10, 4), (7, 8)])
def test_sharded_sampler_equals_distributed_sampler(n, world):
    from torch.utils.data.distributed import DistributedSampler
    ds = list(range(n))
    for rank in range(world):
        for epoch in (0, 3):
            ref = DistributedSampler(ds, num_replicas=world, rank=rank, shuffle=True, seed=0)
            ref.set_epoch(epoch)
            mine = ShardedSampler(n, world, rank, shuffle=True, seed=0)
            mine.set_epoch(epoch)
            assert list(ref) == list(mine)
            assert len(ref) == len(mine) == math.ceil(n / world)


def test_synthetic_dataset_and_transforms():
    ds = SyntheticCIFAR10(64, train=True, transform=TrainTransform())
    x, y = ds[3]
    assert x.shape == (3, 32, 32) and x.dtype == torch.float32 and 0 <= y < 10
    again = SyntheticCIFAR10(64, train=True)
    assert (again.data == ds.data).all() and (again.targets == ds.targets).all()
    imgs = torch.from_numpy(ds.data[:16])
    out = torch.empty(16, 3, 32, 32)
    augment_batch(imgs, out, None, augment=False)
    plain = TrainTransform(augment=False)
    torch.testing.assert_close(out[5], plain(imgs[5]))
    aug = torch.empty(16, 3, 32, 32)
    augment_batch(imgs, aug, torch.Generator().manual_seed(0), augment=True)
    assert torch.isfinite(aug).all() and not torch.equal(aug, out)


def test_batch_pipeline_shards_and_keeps_partial_batch():
    ds = SyntheticCIFAR10(1000, train=True)
    seen = []
    for rank in range(4):
        bp = BatchPipeline(ds, batch=64, world=4, rank=rank, train=True, augment=False, pin=False)
        assert len(bp) == 4  # 250 per rank -> 3 full + partial 58, like the reference (no drop_last)
        sizes = [y.shape[0] for _, y in bp]
        assert sizes == [64, 64, 64, 58]
        seen.append(bp.sampler.indices())
    assert sorted(torch.cat(seen).tolist()) == list(range(1000))


def test_dataloader_parity_settings():
    cfg = TrainConfig(mode="part2b", num_nodes=4, rank=1, train_size=512, test_size=128, num_workers=0)
    train, test = make_loaders(cfg)
    assert train.batch_size == 64 and len(train) == 2 and len(test) == 2
    x, y = next(iter(train))
    assert x.shape == (64, 3, 32, 32) and y.dtype == torch.int64


def test_flat_layout_and_bucket_plans():
    named = list(VGG11().named_parameters())
    L = FlatLayout(named)
    assert L.entries[0].name == "fc1.bias" and L.entries[-1].name == "layers.0.weight"
    assert all(e.offset % 4 == 0 for e in L.entries)
    assert L.total % 1024 == 0 and L.total >= 9_231_114
    layer = L.buckets("layer")
    assert len(layer) == 8 and layer[0].start == 0 and layer[-1].end == L.total
    assert layer[0].names[-1] == "layers.25.weight" and "fc1.weight" in layer[0].names
    for a, b in zip(layer, layer[1:]):
        assert a.end == b.start
    # the reference's DDP probe: 3 buckets of ~9.46 MB / ~27.2 MB / ~0.3 MB
    ddp = L.buckets("ddp")
    assert len(ddp) == 3
    mb = [b.nbytes / 1e6 for b in ddp]
    assert 9.4 < mb[0] < 9.5 and 27.1 < mb[1] < 27.2 and 0.25 < mb[2] < 0.35
    assert len(L.buckets("single")) == 1
    # size-capped plan: closes at conv-layer boundaries, covers everything exactly once, in order
    cap = L.buckets("cap:12")
    assert [round(b.nbytes / 1e6, 1) for b in cap] == [18.9, 14.2, 3.9]
    assert cap[0].start == 0 and cap[-1].end == L.total and all(a.end == b.start for a, b in zip(cap, cap[1:]))
    assert sorted(n for b in cap for n in b.names) == sorted(e.name for e in L.entries)
    assert len(L.buckets("cap:1000")) == 1 and len(L.buckets("cap:0.001")) == 8
    # "+last": the first conv block (the last gradients of backward) in a bucket of its own -- a few KB
    last = L.buckets("cap:16+last")
    assert [round(b.nbytes / 1e6, 1) for b in last[:-1]] == [18.9, 17.7, 0.3] and last[-1].nbytes < 16384
    assert last[-1].names == [e.name for e in L.entries[-4:]] and last[-1].end == L.total
    assert all(a.end == b.start for a, b in zip(last, last[1:]))
    assert sorted(n for b in last for n in b.names) == sorted(e.name for e in L.entries)
    assert len(L.buckets("single+last")) == 2 and len(L.buckets("layer+last")) == 8
    with pytest.raises(ValueError):
        L.buckets("cap:x")
    with pytest.raises(ValueError):
        L.buckets("nonsense")


def test_config_defaults_are_reference_literals_and_validated():
    c = TrainConfig()
    assert (c.lr, c.momentum, c.weight_decay, c.seed, c.global_batch, c.epochs) == (0.1, 0.9, 1e-4, 5000, 256, 1)
    assert TrainConfig(mode="part3", num_nodes=4, rank=0).port == 29508
    assert TrainConfig(mode="part2a", num_nodes=4, rank=3).port == 29501
    assert TrainConfig(mode="part2b", num_nodes=8, rank=0).per_rank_batch == 32
    assert TrainConfig(mode="part2b", num_nodes=8, rank=0, per_rank_batch=256).global_batch == 2048
    with pytest.raises(ValueError):
        TrainConfig(mode="part2b", num_nodes=4, rank=4)
    with pytest.raises(ValueError):
        TrainConfig(mode="part1", num_nodes=2, rank=0)
    with pytest.raises(ValueError):
        TrainConfig(mode="part9")
    # NVLS: automatic for multi-rank runs, never for a single rank, overridable
    assert TrainConfig(mode="part3", num_nodes=8, rank=0).use_nvls
    assert not TrainConfig(mode="part3", num_nodes=1, rank=0).use_nvls
    assert not TrainConfig(mode="part3", num_nodes=8, rank=0, nvls=False).use_nvls
    assert not TrainConfig(mode="part1", nvls=True).use_nvls


def test_real_data_request_without_the_files_is_an_error(tmp_path):
    """A silent fall-back to synthetic images would print CIFAR-looking numbers that are not CIFAR results."""
    import pytest
    from cs744_pytorch_distributed_tutorial_b200.config import TrainConfig
    from cs744_pytorch_distributed_tutorial_b200.data import build_dataset
    cfg = TrainConfig(synthetic=False, data_root=str(tmp_path))
    with pytest.raises(FileNotFoundError, match="cifar-10-batches-py"):
        build_dataset(cfg, True)


def test_batch_pipeline_waits_for_the_copy_of_a_slot_before_refilling_it():
    """The ring of pinned buffers is only safe with a hand-shake: the event behind a batch's host->device copy
    must complete before its buffer is overwritten (the engine passes it through ``consumed``)."""
    ds = SyntheticCIFAR10(64, True, seed=1)
    bp = BatchPipeline(ds, batch=8, train=True, augment=False, pin=False, ring=2)
    log = []

    class Ev:
        def __init__(self, i):
            self.i = i

        def synchronize(self):
            log.append(self.i)

    for i, _ in enumerate(bp):
        bp.consumed(Ev(i))
    # 8 batches through 2 slots: batch i+2 reuses batch i's buffer and must have waited for event i first
    assert log == list(range(6))
