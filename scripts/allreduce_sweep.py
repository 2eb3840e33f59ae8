"""Gradient all-reduce latency / bus bandwidth at the VGG-11 message sizes:
the framework's in-kernel NVLink collectives vs NCCL through torch.distributed.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29700 scripts/allreduce_sweep.py [--blocks 64]

Times are CUDA-event durations over `iters` back-to-back launches, max over ranks.
bus GB/s = 2*(W-1)/W * bytes / time (the NCCL convention)."""
import argparse, json, os, sys
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cs744_pytorch_distributed_tutorial_b200.parallel.comm import Communicator, Store

ap = argparse.ArgumentParser()
ap.add_argument("--blocks", type=int, default=64)
ap.add_argument("--iters", type=int, default=30)
a = ap.parse_args()
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl")
store = Store(os.environ.get("MASTER_ADDR", "127.0.0.1"), int(os.environ["MASTER_PORT"]) + 51, rank, world)
N = 9_232_384
comm = Communicator(store, rank, world, local, heap_bytes=4 * 3 * N + (4 << 20), multicast=True)
grads, params, mom = comm.zeros(N), comm.zeros(N), comm.zeros(N)
grads.normal_()
nccl_buf = torch.randn(N, device="cuda")


def timed(fn):
    for _ in range(5):
        fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.iters):
        fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) * 1e3 / a.iters], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


sizes = [("fc1.bias", 12), ("bn tensor 2 KB", 512), ("bucket 7.7 KB", 1924), ("bucket 0.30 MB", 76_288),
         ("bucket 1.18 MB", 295_936), ("bucket 2.36 MB", 590_848), ("bucket 4.72 MB", 1_180_672),
         ("bucket 9.46 MB", 2_365_452), ("DDP bucket 27.2 MB", 6_789_376), ("all grads 36.9 MB", 9_231_116)]
rows = []
for name, n in sizes:
    n = (n + 3) // 4 * 4
    nbytes = 4 * n
    busf = 2 * (world - 1) / world * nbytes / 1e3   # -> GB/s when divided by us
    r = {"what": name, "bytes": nbytes}
    r["nccl_us"] = timed(lambda: dist.all_reduce(nccl_buf[:n]))
    if nbytes <= 2 << 20:
        r["oneshot_us"] = timed(lambda: comm.allreduce(grads, 0, n, prescale=True, algo=1))
    r["twoshot_us"] = timed(lambda: comm.allreduce(grads, 0, n, prescale=True, algo=2, blocks=a.blocks))
    r["twoshot_sgd_us"] = timed(lambda: comm.allreduce(grads, 0, n, params=params, momentum_buf=mom, lr=0.1, momentum=0.9,
                                                       weight_decay=1e-4, prescale=True, algo=2, blocks=a.blocks))
    r["nvls_us"] = timed(lambda: comm.allreduce(grads, 0, n, algo=3, blocks=a.blocks))
    r["nvls_sgd_us"] = timed(lambda: comm.allreduce(grads, 0, n, params=params, momentum_buf=mom, lr=0.1, momentum=0.9,
                                                    weight_decay=1e-4, algo=3, blocks=a.blocks))
    best = min(v for k, v in r.items() if k in ("oneshot_us", "twoshot_us", "nvls_us") and v)
    r["best_native_us"] = best
    r["native_bus_GBps"] = busf / best
    r["nccl_bus_GBps"] = busf / r["nccl_us"]
    rows.append(r)
    if rank == 0:
        print(json.dumps(r), flush=True)
if rank == 0:
    print(f"# world={world} blocks={a.blocks}")
comm.close(); store.close(); dist.destroy_process_group()
