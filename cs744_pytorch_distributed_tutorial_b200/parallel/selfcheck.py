"""Start-up self-check of every NVLink collective kernel (run by ``bench.py --gpus N`` for N > 1 and by the
multi-GPU tests).

The reference trusts ``torch.distributed`` for its four sync blocks (ref: master/part2a/part2a.py:42-52,
master/part2a/part2a_extra.py:41-58, master/part2b/part2b.py:43-45, DDP at master/part3/part3.py:116); this
framework owns the kernels, so it also owns the proof that they move the right bytes on *this* box:

  * one-shot, two-shot and NVLS all-reduce (mean, both scale orders), the fused all-reduce + SGD, the
    parameter server in pull and push form, broadcast, and the all-gather of owner slices, at the sizes of
    VGG-11's buckets (48 B, 7,696 B, 305,152 B, 9,461,808 B and the whole 36,924,464-byte gradient buffer);
  * a randomised-skew loop (thousands of back-to-back collectives on three sizes with ranks entering late at
    random) that would expose an aliasing epoch flag or a missing fence.

Inputs are small integers, so every sum and every division by a power-of-two world size is exact in fp32:
the expectation is computed on the device with torch ops and compared BIT-EXACTLY, whatever the reduction
order of a kernel.  (World sizes that are not powers of two fall back to a 1e-6 tolerance.)
"""
from __future__ import annotations

import time
from typing import Dict, List, Optional

import torch

from .comm import CH_AUX, CH_MAIN, Communicator, Store

# bytes: fc bias bucket tail, a BatchNorm-sized range, conv 64->128, conv 512->512, the flat gradient buffer
SIZES_BYTES = (48, 7_696, 305_152, 9_461_808, 36_924_464)


def _pattern(rank: int, n: int, salt: int, device) -> torch.Tensor:
    """Integer-valued fp32 data in [-125, 125], different on every rank and for every salt."""
    i = torch.arange(n, device=device, dtype=torch.int64)
    return (((i * 7 + rank * 13 + salt * 31) % 251) - 125).to(torch.float32)


def _mean(world: int, n: int, salt: int, device) -> torch.Tensor:
    acc = torch.zeros(n, device=device)
    for r in range(world):                      # rank order from zero: also the parameter server's order
        acc += _pattern(r, n, salt, device)
    return acc / world


class _Checker:
    def __init__(self, rank: int, exact: bool):
        self.rank, self.exact = rank, exact
        self.cases = 0
        self.failures: List[str] = []

    def check(self, name: str, got: torch.Tensor, want: torch.Tensor) -> None:
        self.cases += 1
        ok = torch.equal(got, want) if self.exact else torch.allclose(got, want, rtol=1e-6, atol=1e-6)
        if not ok:
            bad = int((got != want).sum().item())
            self.failures.append(f"rank {self.rank} {name}: {bad} of {got.numel()} elements differ")


def run_selfcheck(store: Optional[Store], rank: int, world: int, device: int, multicast: bool = True,
                  skew_iters: int = 2000, sizes_bytes=SIZES_BYTES, timeout_s: float = 30.0) -> Dict[str, object]:
    """Collective (every rank calls it).  Returns ``{"ok", "cases", "failures", "seconds", "nvls"}``; the
    caller decides what a failure means (bench.py aborts with a non-zero exit code)."""
    t_begin = time.time()
    dev = torch.device("cuda", device)
    n_max = (max(sizes_bytes) // 4 + 3) // 4 * 4
    heap_bytes = 4 * (3 * n_max + (world - 1) * n_max) + (4 << 20)
    comm = None
    if multicast:
        from ..ops._lib import NativeError
        try:
            comm = Communicator(store, rank, world, device, heap_bytes, timeout_s=timeout_s, multicast=True)
        except NativeError:
            comm = None
    if comm is None:
        comm = Communicator(store, rank, world, device, heap_bytes, timeout_s=timeout_s)
    nvls = bool(comm.has_multicast) and world > 1
    exact = world & (world - 1) == 0
    ck = _Checker(rank, exact)
    grads, params, mom = comm.zeros(n_max), comm.zeros(n_max), comm.zeros(n_max)
    staging = comm.zeros(max((world - 1) * n_max, 4))
    lr, mu, wd = 0.5, 0.5, 0.25                 # powers of two: the fused SGD stays exact on integer data
    salt = 0
    for nbytes in sizes_bytes:
        n = nbytes // 4
        algos = [(1, "oneshot"), (2, "twoshot")] + ([(3, "nvls")] if nvls else [])
        for algo, aname in algos:
            if algo == 1 and n > 256 * 512 * 4:
                continue                        # the one-shot kernel covers one pass of 256 x 512 float4s
            for prescale in ((True, False) if algo != 3 else (True,)):
                salt += 1
                grads[:n].copy_(_pattern(rank, n, salt, dev))
                comm.allreduce(grads, 0, n, prescale=prescale, algo=algo, blocks=32, channel=CH_MAIN)
                ck.check(f"allreduce {aname} {nbytes} B prescale={prescale}", grads[:n], _mean(world, n, salt, dev))
            # fused SGD epilogue: parameters identical on every rank afterwards
            salt += 1
            p0, m0 = _pattern(97, n, salt, dev), _pattern(98, n, salt, dev)
            params[:n].copy_(p0)
            mom[:n].copy_(m0)
            grads[:n].copy_(_pattern(rank, n, salt, dev))
            comm.allreduce(grads, 0, n, params=params, momentum_buf=mom, lr=lr, momentum=mu, weight_decay=wd,
                           prescale=True, algo=algo, blocks=32, channel=CH_MAIN)
            m1 = mu * m0 + (_mean(world, n, salt, dev) + wd * p0)
            ck.check(f"allreduce+sgd {aname} {nbytes} B params", params[:n], p0 - lr * m1)
            if algo != 1:
                # momentum lives on the owner of each slice; the gather completes it on every rank
                comm.allgather_slices(mom, 0, n, channel=CH_MAIN)
                ck.check(f"allgather_slices after {aname} {nbytes} B", mom[:n], m1)
        for push in (False, True):
            salt += 1
            grads[:n].copy_(_pattern(rank, n, salt, dev))
            comm.ps_reduce_bcast(grads, 0, n, root=0, staging=staging if push else None, blocks=32, channel=CH_MAIN)
            ck.check(f"param server push={push} {nbytes} B", grads[:n], _mean(world, n, salt, dev))
        salt += 1
        grads[:n].copy_(_pattern(rank, n, salt, dev))
        comm.broadcast(grads[:n], root=0, channel=CH_AUX)
        ck.check(f"broadcast {nbytes} B", grads[:n], _pattern(0, n, salt, dev))

    # ---- randomised launch skew: the flag protocol must hold when ranks enter thousands of collectives late
    bad = torch.zeros((), device=dev, dtype=torch.int64)
    gen = torch.Generator().manual_seed(99)      # same draws on every rank
    loop_sizes = [s // 4 for s in sizes_bytes[:3]]
    late = torch.randint(0, 4 * world, (skew_iters,), generator=gen).tolist()
    base = [_pattern(rank, n, 7, dev) for n in loop_sizes]
    want = [_mean(world, n, 7, dev) for n in loop_sizes]
    for it in range(skew_iters):
        k = it % len(loop_sizes)
        n = loop_sizes[k]
        grads[:n].copy_(base[k])
        if late[it] == rank:
            torch.cuda._sleep(100_000)           # ~50 us: this rank reaches the entry barrier last
        algo = 1 + it % (3 if nvls else 2)
        comm.allreduce(grads, 0, n, prescale=bool(it & 4), algo=algo, blocks=4, channel=CH_MAIN)
        if exact:
            bad += (grads[:n] != want[k]).sum()
        else:
            bad += ((grads[:n] - want[k]).abs() > 1e-5).sum()
    ck.cases += skew_iters
    n_bad = int(bad.item())
    if n_bad:
        ck.failures.append(f"rank {rank} skew loop: {n_bad} wrong elements over {skew_iters} collectives")
    torch.cuda.synchronize(dev)
    comm.close()
    return {"ok": not ck.failures, "cases": ck.cases, "failures": ck.failures, "nvls": nvls,
            "seconds": round(time.time() - t_begin, 3)}
