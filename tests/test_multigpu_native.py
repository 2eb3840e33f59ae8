"""Multi-GPU tests of the in-kernel NVLink collectives and of the native sync modes (needs >= 2
B200s: ``gpurun --gpus 2 -- python -m pytest tests -m gpu``).  One process per GPU, rendezvous
through the framework's own TCP store, data exchanged only through peer memory."""
import os
import sys
import tempfile

import pytest
import torch
import torch.multiprocessing as mp

sys.path.insert(0, os.path.dirname(__file__))
from helpers import free_port  # noqa: E402

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def _world():
    n = torch.cuda.device_count()
    return 8 if n >= 8 else 4 if n >= 4 else 2


def _rank_data(rank, n, seed=0):
    g = torch.Generator().manual_seed(seed * 1000 + rank)
    return torch.randn(n, generator=g)


def _collectives_worker(rank, world, port, outdir):
    from cs744_pytorch_distributed_tutorial_b200.parallel.comm import CH_AUX, CH_MAIN, Communicator, Store
    torch.cuda.set_device(rank)
    store = Store("127.0.0.1", port, rank, world, timeout_s=60)
    N = 2_400_000  # floats; a conv-layer sized bucket
    comm = Communicator(store, rank, world, rank, heap_bytes=4 * (4 * N + (world - 1) * N) + (1 << 20), timeout_s=30)
    grads, params, mom = comm.zeros(N), comm.zeros(N), comm.zeros(N)
    staging = comm.zeros((world - 1) * N)
    failures = []

    def check(name, got, want, **kw):
        try:
            torch.testing.assert_close(got.cpu(), want, **kw)
        except AssertionError as e:
            failures.append(f"rank {rank} {name}: {str(e)[:300]}")

    # ---- all-reduce: mean of the ranks' buffers, every algorithm / size / scale order
    for case, (start, end) in enumerate([(0, 12), (0, 4100), (4096, 4096 + 65536), (0, N), (1024, N - 2048)]):
        for algo in (1, 2):
            for prescale in (True, False):
                seed = case * 10 + algo * 2 + int(prescale)
                mine = _rank_data(rank, N, seed)
                grads.copy_(mine)
                torch.cuda.synchronize()
                comm.allreduce(grads, start, end, prescale=prescale, algo=algo, blocks=32, channel=CH_MAIN)
                torch.cuda.synchronize()
                mean = sum(_rank_data(r, N, seed) for r in range(world)) / world
                want = mine.clone()
                want[start:end] = mean[start:end]
                check(f"allreduce case{case} algo{algo} prescale{prescale}", grads, want, rtol=1e-5, atol=1e-6)

    # ---- fused SGD epilogue: parameters updated identically on every rank
    for algo in (1, 2):
        start, end = (0, 65536) if algo == 1 else (0, N)
        p0, m0 = _rank_data(99, N, 7), _rank_data(98, N, 7) * 0.1   # same on every rank
        params.copy_(p0); mom.copy_(m0)
        grads.copy_(_rank_data(rank, N, 50 + algo))
        torch.cuda.synchronize()
        comm.allreduce(grads, start, end, params=params, momentum_buf=mom, lr=0.1, momentum=0.9, weight_decay=1e-4,
                       prescale=True, algo=algo, blocks=32)
        torch.cuda.synchronize()
        g = sum(_rank_data(r, N, 50 + algo) / world for r in range(world))
        m1 = 0.9 * m0 + (g + 1e-4 * p0)
        p1 = p0 - 0.1 * m1
        want_p, want_m = p0.clone(), m0.clone()
        want_p[start:end] = p1[start:end]
        check(f"fused sgd params algo{algo}", params, want_p, rtol=1e-5, atol=1e-6)
        if algo == 1:
            want_m[start:end] = m1[start:end]
            check("fused sgd momentum (one-shot: full range on every rank)", mom, want_m, rtol=1e-5, atol=1e-6)
        else:  # two-shot: each rank owns the momentum of its slice only
            n4 = (end - start) // 4
            per = (n4 + world - 1) // world
            lo, hi = start + 4 * min(per * rank, n4), start + 4 * min(per * rank + per, n4)
            check("fused sgd momentum (two-shot: owned slice)", mom[lo:hi], m1[lo:hi], rtol=1e-5, atol=1e-6)

    # ---- parameter server: sum in rank order from zero, then a true division
    for push in (False, True):
        grads.copy_(_rank_data(rank, N, 70 + int(push)))
        torch.cuda.synchronize()
        comm.ps_reduce_bcast(grads, 0, N, root=0, staging=staging if push else None, blocks=32)
        torch.cuda.synchronize()
        total = torch.zeros(N)
        for r in range(world):
            total += _rank_data(r, N, 70 + int(push))
        check(f"param server push={push}", grads, total / world, rtol=0, atol=0)   # bit exact: same order

    # ---- broadcast from rank 0, and a stress loop with launch skew (signal-pad epochs must not alias)
    grads.copy_(_rank_data(rank, N, 80))
    torch.cuda.synchronize()
    comm.broadcast(grads[:4096], root=0, channel=CH_AUX)
    torch.cuda.synchronize()
    check("broadcast", grads[:4096], _rank_data(0, N, 80)[:4096], rtol=0, atol=0)
    ok = True
    small = comm.zeros(4096)
    gen = torch.Generator().manual_seed(5)
    skew = torch.randint(0, 4, (300, world), generator=gen)
    for it in range(300):
        small.fill_(float(rank + it))
        if skew[it, rank] == 0:
            torch.cuda._sleep(200_000)
        comm.allreduce(small, 0, 4096, prescale=False, algo=1 + it % 2, blocks=4, channel=CH_MAIN)
        want = sum(float(r + it) for r in range(world)) / world
        ok = ok and bool((small == want).all().item())
    if not ok:
        failures.append(f"rank {rank}: stress loop produced a wrong all-reduce result")
    torch.cuda.synchronize()
    with open(os.path.join(outdir, f"fail_{rank}.txt"), "w") as f:
        f.write("\n".join(failures))
    comm.close()
    store.close()


def test_collectives_against_host_computed_results():
    world = _world()
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(_collectives_worker, args=(world, free_port(), d), nprocs=world, join=True)
        fails = [open(os.path.join(d, f"fail_{r}.txt")).read() for r in range(world)]
    assert not any(fails), "\n".join(f for f in fails if f)


def _mode_worker(rank, world, mode, fuse, port, outdir):
    from cs744_pytorch_distributed_tutorial_b200.config import TrainConfig
    from cs744_pytorch_distributed_tutorial_b200.engine import Trainer
    torch.cuda.set_device(rank)
    os.environ["LOCAL_RANK"] = str(rank)
    cfg = TrainConfig(mode=mode, master_ip="127.0.0.1", num_nodes=world, rank=rank, port=port, device="cuda",
                      per_rank_batch=32, eval=False, lr=0.01, fuse_optimizer=fuse, timeout_s=60)
    tr = Trainer(cfg)
    nv = tr.backend.native
    init = nv.flat_params.cpu()
    g = torch.Generator().manual_seed(200 + rank)
    losses = []
    for _ in range(2):
        x = torch.randn(32, 3, 32, 32, generator=g).pin_memory()
        y = torch.randint(0, 10, (32,), generator=g).pin_memory()
        losses.append(float(tr.train_step(x, y).item()))
    # DDP semantics: buffers are re-synchronised from rank 0 at the start of the next forward
    tr.backend.sync.resync_buffers()
    torch.cuda.synchronize()
    torch.save({"params": nv.flat_params.cpu(), "init": init, "bn": nv.flat_bn.cpu(), "losses": losses},
               os.path.join(outdir, f"{mode}_{int(fuse)}_{rank}.pt"))
    tr.close()


@pytest.fixture(scope="module")
def mode_results():
    world = _world()
    out = {}
    with tempfile.TemporaryDirectory() as d:
        for mode, fuse in (("part2a", True), ("part2a_extra", True), ("part2b", True), ("part2b", False),
                           ("part3", True), ("part3", False)):
            mp.spawn(_mode_worker, args=(world, mode, fuse, free_port(), d), nprocs=world, join=True)
            out[(mode, fuse)] = [torch.load(os.path.join(d, f"{mode}_{int(fuse)}_{r}.pt")) for r in range(world)]
    return out


def test_replicas_stay_identical_in_every_mode(mode_results):
    for key, ranks in mode_results.items():
        for r in ranks[1:]:
            assert torch.equal(ranks[0]["params"], r["params"]), f"{key}: parameters differ across ranks"


def test_modes_agree_with_each_other(mode_results):
    """Same init, same shards: the first loss is identical and the two-step parameter update points
    the same way in every mode (float atomics + discontinuous ReLU/pool decisions keep it from being
    bit-equal between runs, see DESIGN.md precision policy)."""
    base = mode_results[("part2b", False)][0]
    d0 = base["params"] - base["init"]
    for key, ranks in mode_results.items():
        assert torch.equal(ranks[0]["init"], base["init"])
        # the first loss precedes any exchange; runs differ only through the autotuner's tile choices
        # (accumulation order of bf16 products), i.e. by bf16 rounding noise
        assert abs(ranks[0]["losses"][0] - base["losses"][0]) < 1e-2, key
        d = ranks[0]["params"] - ranks[0]["init"]
        cos = torch.nn.functional.cosine_similarity(d, d0, dim=0).item()
        assert cos > 0.97, f"{key} vs part2b: update cosine {cos}"
        assert abs(d.norm().item() / d0.norm().item() - 1) < 0.05, key


def test_part3_broadcasts_batchnorm_buffers_but_part2b_does_not(mode_results):
    p3 = mode_results[("part3", True)]
    assert all(torch.equal(p3[0]["bn"], r["bn"]) for r in p3[1:])
    p2 = mode_results[("part2b", True)]
    assert not torch.equal(p2[0]["bn"], p2[1]["bn"])      # per-rank statistics


def _nvls_worker(rank, world, port, outdir):
    from cs744_pytorch_distributed_tutorial_b200.parallel.comm import Communicator, Store
    torch.cuda.set_device(rank)
    store = Store("127.0.0.1", port, rank, world, timeout_s=60)
    N = 1_200_000
    comm = Communicator(store, rank, world, rank, heap_bytes=4 * 3 * N + (1 << 20), timeout_s=30, multicast=True)
    failures = []
    if not comm.has_multicast:
        failures.append(f"rank {rank}: multicast mapping missing")
    grads, params, mom = comm.zeros(N), comm.zeros(N), comm.zeros(N)

    def check(name, got, want, **kw):
        try:
            torch.testing.assert_close(got.cpu(), want, **kw)
        except AssertionError as e:
            failures.append(f"rank {rank} {name}: {str(e)[:300]}")

    for case, (start, end) in enumerate([(0, 16), (0, 4100), (1024, N - 2048), (0, N)]):
        mine = _rank_data(rank, N, 300 + case)
        grads.copy_(mine)
        torch.cuda.synchronize()
        comm.allreduce(grads, start, end, algo=3, blocks=32)      # in-switch reduction (multimem)
        torch.cuda.synchronize()
        mean = sum(_rank_data(r, N, 300 + case) for r in range(world)) / world
        want = mine.clone()
        want[start:end] = mean[start:end]
        check(f"nvls allreduce case{case}", grads, want, rtol=1e-5, atol=1e-6)
    # fused SGD through the multicast store of the updated parameters
    p0, m0 = _rank_data(99, N, 8), _rank_data(98, N, 8) * 0.1
    params.copy_(p0); mom.copy_(m0)
    grads.copy_(_rank_data(rank, N, 350))
    torch.cuda.synchronize()
    comm.allreduce(grads, 0, N, params=params, momentum_buf=mom, lr=0.1, momentum=0.9, weight_decay=1e-4, algo=3,
                   blocks=32)
    torch.cuda.synchronize()
    g = sum(_rank_data(r, N, 350) for r in range(world)) / world
    m1 = 0.9 * m0 + (g + 1e-4 * p0)
    check("nvls fused sgd params", params, p0 - 0.1 * m1, rtol=1e-5, atol=1e-6)
    # the peer-mapped (unicast) view of the same VMM heap keeps working: two-shot on it
    grads.copy_(_rank_data(rank, N, 360))
    torch.cuda.synchronize()
    comm.allreduce(grads, 0, N, algo=2, blocks=32)
    torch.cuda.synchronize()
    check("two-shot on the VMM heap", grads, sum(_rank_data(r, N, 360) / world for r in range(world)), rtol=1e-5,
          atol=1e-6)
    with open(os.path.join(outdir, f"fail_{rank}.txt"), "w") as f:
        f.write("\n".join(failures))
    comm.close()
    store.close()


def test_nvls_multimem_allreduce():
    """NVSwitch in-switch reduction (multimem.ld_reduce / multimem.st) over a multicast-mapped heap."""
    world = _world()
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(_nvls_worker, args=(world, free_port(), d), nprocs=world, join=True)
        fails = [open(os.path.join(d, f"fail_{r}.txt")).read() for r in range(world)]
    assert not any(fails), "\n".join(f for f in fails if f)


def _resume_worker(rank, world, mode, phase, port, outdir):
    from cs744_pytorch_distributed_tutorial_b200.config import TrainConfig
    from cs744_pytorch_distributed_tutorial_b200.engine import Trainer
    torch.cuda.set_device(rank)
    os.environ["LOCAL_RANK"] = str(rank)
    ckpt = os.path.join(outdir, f"{mode}.ckpt")
    cfg = TrainConfig(mode=mode, master_ip="127.0.0.1", num_nodes=world, rank=rank, port=port, device="cuda",
                      per_rank_batch=32, eval=False, lr=0.01, timeout_s=60,
                      resume=ckpt if phase == "resume" else None)
    tr = Trainer(cfg)
    nv = tr.backend.native

    def batch(step):
        g = torch.Generator().manual_seed(1000 * step + rank)
        return (torch.randn(32, 3, 32, 32, generator=g).pin_memory(),
                torch.randint(0, 10, (32,), generator=g).pin_memory())

    out = {}
    if phase == "first":
        for s in range(3):
            tr.train_step(*batch(s))
        torch.cuda.synchronize()
        out["mom_before"] = nv.flat_momentum.cpu()
        tr.save_checkpoint(ckpt)             # collective: gathers the owners' momentum slices
        out["mom_after"] = nv.flat_momentum.cpu()
    torch.cuda.synchronize()
    out["p3"] = nv.flat_params.cpu()
    out["mom3"] = nv.flat_momentum.cpu()
    tr.train_step(*batch(3))                 # the step both runs take from the same state
    torch.cuda.synchronize()
    out["p4"] = nv.flat_params.cpu()
    torch.save(out, os.path.join(outdir, f"{mode}_{phase}_{rank}.pt"))
    tr.close()


@pytest.mark.parametrize("mode", ["part3", "part2b"])
def test_checkpoint_holds_the_full_momentum_and_resume_continues_the_run(mode):
    """Fused modes keep SGD momentum on the owner of each 1/W slice; the checkpoint must still hold
    torch.optim.SGD's full state (ref: optimizer at master/part3/part3.py:117-118) and a run resumed from
    it in a fresh process group must take the step the uninterrupted run takes."""
    world = _world()
    with tempfile.TemporaryDirectory() as d:
        res = {}
        for phase in ("first", "resume"):
            mp.spawn(_resume_worker, args=(world, mode, phase, free_port(), d), nprocs=world, join=True)
            res[phase] = [torch.load(os.path.join(d, f"{mode}_{phase}_{r}.pt")) for r in range(world)]
        ck = torch.load(os.path.join(d, f"{mode}.ckpt"), weights_only=False)
    first, resumed = res["first"], res["resume"]
    # before the gather a non-owner's copy differs from the owner's; afterwards every rank holds the same
    assert not torch.equal(first[0]["mom_before"], first[1]["mom_before"])
    for r in first[1:]:
        assert torch.equal(first[0]["mom_after"], r["mom_after"]), "momentum differs across ranks after the gather"
    # every rank's live values survived where it is the owner
    for r, rec in enumerate(first):
        same = (rec["mom_before"] == rec["mom_after"]).float().mean().item()
        assert same > 1.0 / world - 0.05, f"rank {r}: the gather overwrote its own slices ({same:.2f} unchanged)"
    # no stale zero slices: the big conv tensors must be dense in the saved state
    for name, v in ck["optimizer"]["momentum_buffers"].items():
        if v.dim() == 4 and v.numel() > 100_000:
            frac = (v != 0).float().mean().item()
            assert frac > 0.95, f"{name}: only {frac:.2f} of the saved momentum is non-zero"
    # the fresh process group starts from exactly the saved state, on every rank ...
    for r in range(world):
        assert torch.equal(resumed[r]["p3"], first[0]["p3"]), "the resumed run did not start from the saved parameters"
        assert torch.equal(resumed[r]["mom3"], first[0]["mom_after"]), "the resumed run lost optimizer state"
    # ... and takes the step the uninterrupted run takes (same state, same batch; the two differ only by the
    # accumulation order of bf16 products and float atomics).  Momentum carries ~2.7x the gradient after three
    # steps, so a lost slice would show up as a step of the wrong length and direction.
    d_full = first[0]["p4"] - first[0]["p3"]
    d_res = resumed[0]["p4"] - resumed[0]["p3"]
    cos = torch.nn.functional.cosine_similarity(d_full, d_res, dim=0).item()
    ratio = d_res.norm().item() / d_full.norm().item()
    assert cos > 0.98, f"{mode}: resumed step cosine {cos}"
    assert abs(ratio - 1) < 0.05, f"{mode}: resumed step length ratio {ratio}"


def _selfcheck_worker(rank, world, port, outdir):
    import json
    from cs744_pytorch_distributed_tutorial_b200.parallel.comm import Store
    from cs744_pytorch_distributed_tutorial_b200.parallel.selfcheck import run_selfcheck
    torch.cuda.set_device(rank)
    store = Store("127.0.0.1", port, rank, world, timeout_s=60)
    res = run_selfcheck(store, rank, world, rank, multicast=True, skew_iters=2000)
    with open(os.path.join(outdir, f"selfcheck_{rank}.json"), "w") as f:
        json.dump(res, f)
    store.close()


def test_collective_selfcheck_is_bit_exact_at_the_bucket_sizes_and_under_launch_skew():
    """The start-up check bench.py runs at N > 1: every collective kernel at VGG-11's bucket sizes against a
    device-computed expectation on integer data (bit-exact for power-of-two worlds) and 2,000 skewed launches."""
    import json
    world = _world()
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(_selfcheck_worker, args=(world, free_port(), d), nprocs=world, join=True)
        res = [json.load(open(os.path.join(d, f"selfcheck_{r}.json"))) for r in range(world)]
    for r in res:
        assert r["ok"], r["failures"][:5]
        assert r["cases"] > 2000
    print("selfcheck seconds per rank:", [r["seconds"] for r in res], "nvls:", res[0]["nvls"])


def _oracle_worker(rank, world, mode, port, outdir):
    from cs744_pytorch_distributed_tutorial_b200.config import TrainConfig
    from cs744_pytorch_distributed_tutorial_b200.engine import Trainer
    torch.cuda.set_device(rank)
    os.environ["LOCAL_RANK"] = str(rank)
    cfg = TrainConfig(mode=mode, master_ip="127.0.0.1", num_nodes=world, rank=rank, port=port, device="cuda",
                      per_rank_batch=32, eval=False, lr=0.01, fuse_optimizer=False, cuda_graph=False, timeout_s=60)
    tr = Trainer(cfg)
    be, nv = tr.backend, tr.backend.native
    g = torch.Generator().manual_seed(400 + rank)
    x = torch.randn(32, 3, 32, 32, generator=g).cuda()
    y = torch.randint(0, 10, (32,), generator=g).cuda()
    # local gradient of this rank's shard: backward with the bucket hook detached, so nothing is exchanged yet
    hook, nv.on_bucket_ready = nv.on_bucket_ready, None
    be.sync.pre_forward()
    nv.forward(x, training=True)
    be.sync.post_forward()
    nv.loss_fwd_bwd(y)
    nv.backward()
    torch.cuda.synchronize()
    local = nv.flat_grads.cpu()
    nv.on_bucket_ready = hook
    if mode == "part3":
        for k in range(len(be.sync.buckets)):       # what backward() would have queued, on the same gradients
            be.sync.reducer.bucket_ready(k)
    be.sync.sync_only()
    torch.cuda.synchronize()
    torch.save({"local": local, "synced": nv.flat_grads.cpu()}, os.path.join(outdir, f"oracle_{mode}_{rank}.pt"))
    tr.close()


@pytest.mark.parametrize("mode", ["part2a", "part2a_extra", "part2b", "part3"])
def test_synced_gradient_is_the_mean_of_the_per_rank_gradients(mode):
    """The GPU-side mode oracle: with the optimizer un-fused, what a sync mode leaves in ``flat_grads`` on every
    rank is the mean of the W per-rank gradients the SAME executor computed -- bit-exact for the parameter
    server's rank-order sum and true division (ref: master/part2a/part2a.py:42-52, part2a_extra.py:41-58),
    to fp32 rounding for the divide-then-sum orders (ref: master/part2b/part2b.py:43-45, DDP at
    master/part3/part3.py:116)."""
    world = _world()
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(_oracle_worker, args=(world, mode, free_port(), d), nprocs=world, join=True)
        res = [torch.load(os.path.join(d, f"oracle_{mode}_{r}.pt")) for r in range(world)]
    for r in res[1:]:
        assert torch.equal(res[0]["synced"], r["synced"]), f"{mode}: ranks hold different synced gradients"
    if mode in ("part2a", "part2a_extra"):
        acc = torch.zeros_like(res[0]["local"])
        for r in res:
            acc += r["local"]
        assert torch.equal(res[0]["synced"], acc / world)
    else:
        want = sum(r["local"] / world for r in res)
        # summation order differs (switch / peer order): 1e-6 of the largest addend of the tensor
        atol = 1e-6 * max(float(r["local"].abs().max()) for r in res)
        torch.testing.assert_close(res[0]["synced"], want, rtol=1e-6, atol=atol)
