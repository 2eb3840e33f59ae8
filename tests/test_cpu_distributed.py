"""CPU / gloo multi-process tests of the plumbing path (world_size 2 and 4): every sync mode must
produce the *mean gradient over ranks* and the same one-step parameters as a single process that
emulates the W ranks (per-rank BatchNorm statistics)."""
import os
import sys
import tempfile

import pytest
import torch
import torch.multiprocessing as mp
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(__file__))
from helpers import free_port  # noqa: E402

from cs744_pytorch_distributed_tutorial_b200.config import TrainConfig  # noqa: E402
from cs744_pytorch_distributed_tutorial_b200.models.vgg import VGG11  # noqa: E402

BATCH = 4  # per rank; tiny so that VGG-11 on CPU stays fast


def _rank_batch(rank):
    g = torch.Generator().manual_seed(100 + rank)
    return torch.randn(BATCH, 3, 32, 32, generator=g), torch.randint(0, 10, (BATCH,), generator=g)


def _worker(rank, world, mode, port, outdir):
    from cs744_pytorch_distributed_tutorial_b200.engine import Trainer
    torch.set_num_threads(1)
    cfg = TrainConfig(mode=mode, master_ip="127.0.0.1", num_nodes=world, rank=rank, port=port, device="cpu",
                      per_rank_batch=BATCH, cpu_threads=1, eval=False, timeout_s=60)
    tr = Trainer(cfg)
    x, y = _rank_batch(rank)
    loss = tr.train_step(x, y)
    out = {"loss": float(loss), "params": {k: v.detach().clone() for k, v in tr.model.named_parameters()},
           "grads": {k: v.grad.detach().clone() for k, v in tr.model.named_parameters()},
           "buffers": {k: v.detach().clone() for k, v in tr.model.named_buffers()}}
    if mode == "part3":
        out["launch_log"] = list(tr.backend.sync.launch_log)
    torch.save(out, os.path.join(outdir, f"{mode}_{rank}.pt"))
    tr.close()


def _run(mode, world):
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(_worker, args=(world, mode, free_port(), d), nprocs=world, join=True)
        return [torch.load(os.path.join(d, f"{mode}_{r}.pt"), weights_only=False) for r in range(world)]


def _emulated(world):
    """One process: W replicas' gradients averaged in rank order, one SGD step (torch reference)."""
    torch.manual_seed(5000)
    model = VGG11()
    opt = torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    grads = None
    init = {k: v.detach().clone() for k, v in model.state_dict().items()}
    for r in range(world):
        model.load_state_dict(init)  # every rank starts from the same weights and buffers
        model.train()
        model.zero_grad()
        x, y = _rank_batch(r)
        F.cross_entropy(model(x), y).backward()
        g = [p.grad.detach().clone() for p in model.parameters()]
        grads = g if grads is None else [a + b for a, b in zip(grads, g)]
    model.load_state_dict(init)
    for p, g in zip(model.parameters(), grads):
        p.grad = g / world
    opt.step()
    return {k: v.detach().clone() for k, v in model.named_parameters()}, \
           {k: g / world for (k, _), g in zip(model.named_parameters(), grads)}


@pytest.fixture(scope="module")
def results_w2():
    return {m: _run(m, 2) for m in ("part2a", "part2a_extra", "part2b", "part3")}


def test_all_modes_agree_with_emulated_run_world2(results_w2):
    params_ref, grads_ref = _emulated(2)
    for mode, ranks in results_w2.items():
        for r, out in enumerate(ranks):
            for k in grads_ref:
                torch.testing.assert_close(out["grads"][k], grads_ref[k], rtol=1e-4, atol=1e-6,
                                           msg=lambda m, mode=mode, k=k: f"{mode} rank {r} grad {k}: {m}")
                torch.testing.assert_close(out["params"][k], params_ref[k], rtol=1e-4, atol=1e-6)


def test_replicas_identical_after_step(results_w2):
    for mode, ranks in results_w2.items():
        for k in ranks[0]["params"]:
            assert torch.equal(ranks[0]["params"][k], ranks[1]["params"][k]), f"{mode}: {k} differs across ranks"


def test_param_server_variants_bit_identical(results_w2):
    """gather/scatter and isend/irecv compute the same sum in the same order."""
    a, b = results_w2["part2a"][0], results_w2["part2a_extra"][0]
    for k in a["grads"]:
        assert torch.equal(a["grads"][k], b["grads"][k]), k


def test_ddp_mode_syncs_buffers_and_launches_buckets_in_order(results_w2):
    r0, r1 = results_w2["part3"]
    assert r0["launch_log"] == list(range(4)) == r1["launch_log"]      # default plan cap:12+last -> four buckets on VGG-11
    # part2b keeps per-rank BatchNorm statistics ...
    b0, b1 = results_w2["part2b"]
    assert not torch.equal(b0["buffers"]["layers.1.running_mean"], b1["buffers"]["layers.1.running_mean"])


def test_world4_part2b_matches_emulation():
    ranks = _run("part2b", 4)
    params_ref, grads_ref = _emulated(4)
    for out in ranks:
        for k in grads_ref:
            torch.testing.assert_close(out["grads"][k], grads_ref[k], rtol=1e-4, atol=1e-6)
            torch.testing.assert_close(out["params"][k], params_ref[k], rtol=1e-4, atol=1e-6)
