"""Constructor paths of the GPU sync strategies, executed on CPU against stand-ins.

parallel/sync_native.py only runs for real on a GPU; a typo in a constructor would take every part3 run
down with it.  These tests build the strategies around a fake executor / communicator / reducer (no CUDA
call is reached in a constructor) for the default configuration and for each experimental toggle, and
check the bookkeeping they derive: bucket plan hand-off, two-phase split, staging layout of the fused
wgrad -> reduce-scatter path.
"""
import types

import pytest
import torch

from cs744_pytorch_distributed_tutorial_b200.models.native_vgg import FlatLayout
from cs744_pytorch_distributed_tutorial_b200.models.vgg import VGG11
from cs744_pytorch_distributed_tutorial_b200.parallel import sync_native


class FakeReducer:
    def __init__(self, comm, buckets, *a, **kw):
        self.n_buckets, self.kw, self.staged, self.ready = len(buckets), kw, None, []

    def bucket_ready(self, k):
        self.ready.append(k)

    def set_staged(self, stage_off, lo, hi, base):
        self.staged = (stage_off, list(lo), list(hi), list(base))

    def set_lazy_exit(self, on=True):
        self.lazy_exit = on

    def stream(self, device):
        return "reducer-stream"

    def finalize(self):
        pass

    def close(self):
        pass


class FakeComm:
    def __init__(self, world, rank=0):
        self.world, self.rank, self._cursor = world, rank, 0

    def zeros(self, numel):
        t = torch.zeros(numel)
        t._off = self._cursor
        self._cursor += 4 * numel
        return t

    def offset_of(self, t):
        return getattr(t, "_off", 0)

    def peer_ptr(self, r):
        return 0x10000000 * (r + 1)


def fake_native():
    model = VGG11()
    layout = FlatLayout(list(model.named_parameters()))
    nv = types.SimpleNamespace(layout=layout, device=torch.device("cpu"), on_bucket_ready=None, wgrad_push=None)
    nv.blocks = []
    for i, s in enumerate(model.block_specs()):
        nv.blocks.append(types.SimpleNamespace(spec=s, first=(i == 0), w_master=None, wf=None, wd=None,
                                               taps=1 if i == 0 else 9, kin=64 if i == 0 else s.cin,
                                               prep=(None, None, None, s.cout, s.cin, 9, 64 if i == 0 else 9 * s.cin)))
    nv.flat_grads = nv.flat_params = nv.flat_momentum = torch.zeros(4)

    def set_bucket_plan(buckets):
        nv.buckets = list(buckets)
        nv._bucket_of_block = {}
        for bk in buckets:
            idxs = [int(n.split(".")[1]) for n in bk.names if n.startswith("layers.")]
            blocks = [j for j, b in enumerate(nv.blocks) if any(x in (b.spec.conv_idx, b.spec.bn_idx) for x in idxs)]
            nv._bucket_of_block.setdefault(min(blocks) if blocks else len(nv.blocks) - 1, []).append(bk.index)
    nv.set_bucket_plan = set_bucket_plan
    nv.blocks_of_bucket = lambda k: [i for i, ks in nv._bucket_of_block.items() if k in ks]
    nv._dgrad_done = [object() for _ in nv.blocks]
    return nv


@pytest.fixture(autouse=True)
def _stubs(monkeypatch):
    monkeypatch.setattr(sync_native, "BucketReducer", FakeReducer)
    monkeypatch.setattr(sync_native.K, "BatchedWeightPrep", lambda specs: (lambda: None))
    monkeypatch.delenv("B200DP_WGRAD_PUSH", raising=False)

    class _Fake:                      # LocalSGD forks its first optimizer phase onto a stream of its own
        def __init__(self, *a, **kw):
            pass

        def record(self, *a):
            pass

        def wait_event(self, *a):
            pass
    import contextlib
    monkeypatch.setattr(torch.cuda, "Stream", _Fake)
    monkeypatch.setattr(torch.cuda, "Event", _Fake)
    monkeypatch.setattr(torch.cuda, "stream", lambda s: contextlib.nullcontext())
    monkeypatch.setattr(torch.cuda, "current_stream", lambda *a: _Fake())


def _make(world, **kw):
    nv = fake_native()
    s = sync_native.BucketedOverlapSync(nv, FakeComm(world), lr=0.1, momentum=0.9, weight_decay=1e-4,
                                        fuse_optimizer=True, **kw)
    return nv, s


def test_default_part3_strategy_constructs_and_hands_buckets_to_the_reducer():
    nv, s = _make(2, bucket_plan="layer", blocks=32)
    assert len(s.buckets) == 8 and s.reducer.kw["blocks"] == 32
    assert not s._wgrad_push and not s.handles_shadows      # the engine re-lays the operand copies out after the step
    assert nv.wgrad_push is None and s.reducer.staged is None
    nv.on_bucket_ready(3)
    assert s.reducer.ready == [3]
    s.sync_and_step()
    s.close()
    assert nv.on_bucket_ready is None
    for plan in ("ddp", "single", "cap:12", "cap:16+last"):
        _make(4, bucket_plan=plan)
    assert [len(b.names) for b in _make(4)[1].buckets] == [10, 8, 12, 4]     # default plan cap:12+last


@pytest.mark.parametrize("world", [2, 4, 8])
def test_wgrad_push_toggle_lays_out_staging_consistently(monkeypatch, world):
    monkeypatch.setenv("B200DP_WGRAD_PUSH", "1")
    nv, s = _make(world, bucket_plan="layer")
    assert s._wgrad_push and sorted(nv.wgrad_push) == list(range(1, len(nv.blocks)))   # layer 0 stays on the all-reduce
    stage_off, lo, hi, base = s.reducer.staged
    end_of_stage = 0
    for k, bk in enumerate(s.buckets):
        n4 = (bk.end - bk.start) // 4
        per4 = (n4 + world - 1) // world
        assert base[k] == end_of_stage and base[k] % 4 == 0          # areas are disjoint and packed
        end_of_stage += world * per4 * 4
        assert bk.start <= lo[k] <= hi[k] <= bk.end and lo[k] % 4 == 0 and hi[k] % 4 == 0
    assert s._stage.numel() >= end_of_stage
    for i, cfg in nv.wgrad_push.items():
        blk = nv.blocks[i]
        k = next(j for j, b in enumerate(s.buckets) if b.start <= cfg["flat_off"] < b.end)
        assert (cfg["own_start"], cfg["own_len"]) == (s.buckets[k].start, s.buckets[k].end - s.buckets[k].start)
        assert cfg["stage_base"] == base[k] and (lo[k], hi[k]) == (cfg["flat_off"],
                                                                   cfg["flat_off"] + blk.spec.cout * 9 * blk.spec.cin)
        assert cfg["counters"].numel() >= 9 * ((blk.spec.cout + 127) // 128) * (blk.spec.cin // 64)
        assert cfg["world"] == world and len(cfg["peer_stage"]) == world
    # the first bucket also holds the classifier: its pushed range must not cover it
    fc = nv.layout.by_name["fc1.weight"]
    assert not (lo[0] <= fc.offset < hi[0])


def test_make_native_sync_dispatches_every_mode(monkeypatch):
    from cs744_pytorch_distributed_tutorial_b200.config import TrainConfig
    kinds = {}
    for mode, world in (("part1", 1), ("part3", 1), ("part2a", 4), ("part2a_extra", 4), ("part2b", 4), ("part3", 4)):
        nv = fake_native()
        comm = FakeComm(world) if world > 1 else None
        cfg = TrainConfig(mode=mode, num_nodes=world, rank=0, device="cpu")
        s = sync_native.make_native_sync(mode, nv, comm, cfg)
        kinds[(mode, world)] = type(s).__name__
        s.close()
    assert kinds == {("part1", 1): "LocalSGD", ("part3", 1): "LocalSGD", ("part2a", 4): "ParamServerSync",
                     ("part2a_extra", 4): "ParamServerSync", ("part2b", 4): "AllReduceSync",
                     ("part3", 4): "BucketedOverlapSync"}
    # one rank but the collective path kept alive (tests, profiling)
    nv = fake_native()
    cfg = TrainConfig(mode="part3", num_nodes=1, rank=0, device="cpu", single_rank_collectives=True)
    s = sync_native.make_native_sync("part3", nv, FakeComm(1), cfg)
    assert type(s).__name__ == "BucketedOverlapSync" and s.reducer.kw["blocks"] == cfg.comm_blocks == 32


def test_local_sgd_two_phase_split_covers_all_parameters():
    nv = fake_native()
    nv._wgrad_stream = None
    s = sync_native.LocalSGD(nv, None, lr=0.1, momentum=0.9, weight_decay=1e-4)
    if s.overlap:
        assert 0 < s.split_end < nv.layout.total and s.handles_shadows
        assert s.split_end == s.buckets[s.split_bucket].end
        big = sum(b.end - b.start for b in s.buckets[: s.split_bucket + 1])
        assert big / nv.layout.total > 0.9          # phase 1 carries the bulk of the bytes


def test_sync_impl_option_selects_the_allreduce_kernel():
    from cs744_pytorch_distributed_tutorial_b200.config import TrainConfig
    for name, code in (("auto", 0), ("oneshot", 1), ("twoshot", 2), ("nvls", 3)):
        cfg = TrainConfig(mode="part2b", num_nodes=2, rank=0, device="cpu", sync_impl=name)
        s = sync_native.make_native_sync("part2b", fake_native(), FakeComm(2), cfg)
        assert type(s).__name__ == "AllReduceSync" and s.algo == code
    with pytest.raises(ValueError):
        TrainConfig(mode="part2b", num_nodes=2, rank=0, sync_impl="ring")
