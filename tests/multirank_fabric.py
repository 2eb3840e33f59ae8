"""In-process stand-in for the NVLink side of part3, for tests/test_cpu_multirank_emulated.py.

W executors (one per "rank") live in one process.  Their communicator, bucket reducer and the push
flavour of the weight-gradient kernel are replaced by objects that talk to a shared `Fabric`:

  * `FabricComm`      -- symmetric-heap allocation and broadcast (copies the root executor's tensor);
  * `FabricReducer`   -- records bucket_ready(k); the collectives run when `Fabric.run_collectives()` is
                         called after every rank has finished backward (legal: on the GPU each bucket kernel
                         is ordered after the wgrad that completes the bucket, on every rank);
  * `push_tile_handoff` -- what conv_wgrad_push_kernel does after the GEMM: every float4 of the finished
                         tensor that another rank owns goes into that rank's staging slot.

The all-reduce + SGD and the staged reduce + SGD below use the SAME index arithmetic as
allreduce_twoshot_kernel / reduce_staged_sgd_kernel (csrc/collectives.cu): per-bucket float4 slices of
ceil(n4 / W), staging slot of source rank s at stage_base + s * per4 * 4.
"""
import torch

PEER_STRIDE = 1 << 40


class Fabric:
    def __init__(self, world):
        self.world = world
        self.execs = [None] * world      # NativeVGG per rank
        self.syncs = [None] * world      # BucketedOverlapSync per rank
        self.reducers = [None] * world
        self.done = {}                   # post-backward collectives already performed, per kind

    # ---- collectives, executed once every rank has queued its buckets ---------------------------------
    def run_collectives(self):
        W = self.world
        ready = [r.ready for r in self.reducers]
        assert all(x == ready[0] for x in ready), "ranks queued different bucket sequences"
        for k in ready[0]:
            red0 = self.reducers[0]
            bk = red0.buckets[k]
            n4 = (bk.end - bk.start) // 4
            per4 = (n4 + W - 1) // W
            staged = red0.staged is not None and red0.staged[2][k] > red0.staged[1][k]
            new_params = {}
            for r in range(W):
                lo4 = min(per4 * r, n4)
                hi4 = min(lo4 + per4, n4)
                if hi4 <= lo4:
                    continue
                e_lo, e_hi = bk.start + 4 * lo4, bk.start + 4 * hi4
                if staged:
                    g = self._staged_sum(r, k, bk, per4, lo4, e_lo, e_hi)
                else:
                    g = sum(self.execs[s].flat_grads[e_lo:e_hi] / W for s in range(W))      # pre-divide, then sum
                    g = g.clone()
                if staged:
                    g = g / W
                red = self.reducers[r]
                if not red.fuse:
                    for s in range(W):                            # all-gather of the mean gradient
                        self.execs[s].flat_grads[e_lo:e_hi] = g
                    continue
                p = self.execs[r].flat_params[e_lo:e_hi]
                m = self.execs[r].flat_momentum[e_lo:e_hi]
                m.mul_(red.momentum).add_(g + red.weight_decay * p)
                new_params[(e_lo, e_hi)] = p - red.lr * m
                self.execs[r].flat_grads[e_lo:e_hi] = g          # the owner keeps the mean gradient
            for (e_lo, e_hi), v in new_params.items():           # all-gather of the updated parameters
                for s in range(W):
                    self.execs[s].flat_params[e_lo:e_hi] = v
        for r in self.reducers:
            r.ready = []

    def _staged_sum(self, r, k, bk, per4, lo4, e_lo, e_hi):
        """reduce_staged_sgd_kernel, reduction part, for owner r."""
        W = self.world
        stage_off, pushed_lo, pushed_hi, stage_base = self.reducers[r].staged
        stage = self.syncs[r]._stage
        g = torch.zeros(e_hi - e_lo)
        p_lo, p_hi = max(pushed_lo[k], e_lo), min(pushed_hi[k], e_hi)
        # outside the pushed range (biases, BatchNorm, classifier): pull from every rank
        outside = [(e_lo, p_lo), (p_hi, e_hi)] if p_hi > p_lo else [(e_lo, e_hi)]
        for a, b in outside:
            if b > a:
                g[a - e_lo: b - e_lo] = sum(self.execs[s].flat_grads[a:b] for s in range(W))
        if p_hi > p_lo:
            acc = self.execs[r].flat_grads[p_lo:p_hi].clone()
            first4 = (p_lo - bk.start) // 4 - lo4              # index inside my slice
            for s in range(W):
                if s == r:
                    continue
                off = stage_base[k] + (s * per4 + first4) * 4
                acc += stage[off: off + (p_hi - p_lo)]
            g[p_lo - e_lo: p_hi - e_lo] = acc
        return g


class FabricComm:
    def __init__(self, fabric, rank):
        self.fabric, self.rank, self.world = fabric, rank, fabric.world
        self._calls = {}

    def zeros(self, numel, dtype=torch.float32, align=256):
        return torch.zeros(numel, dtype=dtype)

    def offset_of(self, t):
        return 4096

    def peer_ptr(self, r):
        return (r + 1) * PEER_STRIDE

    def prefer_nvls(self, on=True):
        pass

    def broadcast(self, t, root=0, blocks=0, channel=0, stream=None):
        mine, src = self.fabric.execs[self.rank], self.fabric.execs[root]
        if mine is None or src is None or mine is src:
            return
        for name in ("flat_params", "bn_state"):
            if getattr(mine, name).data_ptr() == t.data_ptr():
                t.copy_(getattr(src, name))
                return
        raise AssertionError("broadcast of a tensor the fabric does not know")

    # ---- post-backward collectives (part2a / part2a_extra / part2b): the first rank to call performs the
    # ---- collective for everybody (all gradients exist by then), the later calls of the same round are no-ops
    def _once(self, what):
        n = self._calls.get(what, 0) + 1
        self._calls[what] = n
        done = self.fabric.done.get(what, 0)
        if done >= n:
            return False
        self.fabric.done[what] = n
        return True

    def allreduce(self, grads, start, end, *, params=None, momentum_buf=None, lr=0.0, momentum=0.0, weight_decay=0.0,
                  prescale=True, algo=0, blocks=0, channel=0, stream=None):
        if not self._once("allreduce"):
            return
        ex, W = self.fabric.execs, self.world
        if prescale:
            mean = sum(e.flat_grads[start:end] / W for e in ex)
        else:
            mean = sum(e.flat_grads[start:end] for e in ex) / W
        if params is None:
            for e in ex:
                e.flat_grads[start:end] = mean
            return
        n4 = (end - start) // 4
        if n4 * 16 <= 256 * 1024:
            # one-shot kernel: every rank reduces the whole range and updates its own momentum and parameters
            for e in ex:
                e.flat_grads[start:end] = mean
                p, m = e.flat_params[start:end], e.flat_momentum[start:end]
                m.mul_(momentum).add_(mean + weight_decay * p)
                p.sub_(lr * m)
            return
        # two-shot / NVLS kernels: momentum and the mean gradient exist only on the OWNER of each 1/W slice
        per4 = (n4 + W - 1) // W
        for r in range(W):
            lo4 = min(per4 * r, n4)
            hi4 = min(lo4 + per4, n4)
            if hi4 <= lo4:
                continue
            a, b = start + 4 * lo4, start + 4 * hi4
            g = mean[a - start: b - start]
            p, m = ex[r].flat_params[a:b], ex[r].flat_momentum[a:b]
            m.mul_(momentum).add_(g + weight_decay * p)
            new_p = p - lr * m
            ex[r].flat_grads[a:b] = g
            for e in ex:
                e.flat_params[a:b] = new_p

    def allgather_slices(self, t, start, end, blocks=0, channel=0, stream=None):
        """allgather_slices_kernel: every owner's slice of `flat_momentum[start:end]` goes to every rank."""
        if not self._once(("gather", start, end)):
            return
        ex, W = self.fabric.execs, self.world
        n4 = (end - start) // 4
        per4 = (n4 + W - 1) // W
        for r in range(W):
            lo4 = min(per4 * r, n4)
            hi4 = min(lo4 + per4, n4)
            a, b = start + 4 * lo4, start + 4 * hi4
            for e in ex:
                if e is not ex[r]:
                    e.flat_momentum[a:b] = ex[r].flat_momentum[a:b]

    def ps_reduce_bcast(self, grads, start, end, *, root=0, staging=None, blocks=0, channel=0, stream=None):
        if not self._once("ps"):
            return
        ex, W = self.fabric.execs, self.world
        acc = torch.zeros(end - start)
        for e in ex:                                   # rank order, starting from zero
            acc = acc + e.flat_grads[start:end]
        mean = acc / W
        for e in ex:
            e.flat_grads[start:end] = mean

    def close(self):
        pass


class FabricReducer:
    def __init__(self, comm, buckets, grads, params, momentum_buf, lr, momentum, weight_decay, prescale=True, blocks=16,
                 channel=1):
        self.comm, self.buckets = comm, list(buckets)
        self.lr, self.momentum, self.weight_decay = lr, momentum, weight_decay
        self.n_buckets, self.ready, self.staged = len(self.buckets), [], None
        self.fuse = params is not None          # False: all-reduce only, the strategy runs SGD itself
        comm.fabric.reducers[comm.rank] = self

    def bucket_ready(self, k):
        self.ready.append(k)

    def set_staged(self, stage_off, lo, hi, base):
        self.staged = (stage_off, list(lo), list(hi), list(base))

    def finalize(self):
        pass

    def set_lazy_exit(self, on=True):
        pass

    def stream(self, device):
        class _Stream:
            def wait_event(self, ev):
                pass
        return _Stream()

    def close(self):
        pass


def push_tile_handoff(fabric, dw_flat, *, world, rank, peer_stage, flat_off, own_start, own_len, stage_base):
    """conv_wgrad_push_kernel's tile hand-off for a finished tensor `dw_flat` (its float4s, all tiles)."""
    n4 = own_len // 4
    per4 = (n4 + world - 1) // world
    first4 = (flat_off - own_start) // 4
    count4 = dw_flat.numel() // 4
    for owner in range(world):
        if owner == rank:
            continue
        a4 = max(first4, owner * per4)
        b4 = min(first4 + count4, (owner + 1) * per4, n4)
        if b4 <= a4:
            continue
        decoded = int(peer_stage[owner]) // PEER_STRIDE - 1      # the pointer the kernel would store through
        assert decoded == owner
        stage = fabric.syncs[owner]._stage
        dst = stage_base + (rank * per4 + (a4 - owner * per4)) * 4
        src = (a4 - first4) * 4
        stage[dst: dst + (b4 - a4) * 4] = dw_flat[src: src + (b4 - a4) * 4]
