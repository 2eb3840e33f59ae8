"""sm_100a executor for the VGG family: flat parameter/gradient/momentum buffers and a
static forward/backward kernel schedule.

Design:
  * All parameters live in ONE flat fp32 buffer, gradients in a second, SGD momentum in
    a third.  ``module.weight.data`` / ``.grad`` are *views* into them, so the kernels
    write gradients straight into (optionally peer-visible) memory and the collectives
    and the fused optimizer work on contiguous ranges -- no copy in or out.
  * The flat order is the *gradient-ready* order (reverse definition order), so that every
    bucket of the overlap reducer is one contiguous range.
  * Conv weights are stored KRSC ``[Cout][3][3][Cin]``; the module sees them as OIHW
    tensors with channels_last strides, which keeps ``state_dict`` compatibility
     without a layout conversion in the hot path.
  * Activations are bf16 NHWC in pre-allocated workspaces; the step allocates nothing and
    never synchronises with the host, so it can be captured in a CUDA graph.

The kernel schedule replaces what the reference gets from ``model(data)`` /
``loss.backward()`` (ref: master/part1/part1.py:35-37).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn as nn

from ..ops import native as K

# weight gradients on a second stream, concurrently with the dgrad / BatchNorm-backward chain

ALIGN = 4  # floats; every tensor starts on a 16-byte boundary inside the flat buffers


@dataclass
class FlatEntry:
    name: str
    offset: int      # in floats
    numel: int
    shape: Tuple[int, ...]      # logical (state_dict) shape
    kind: str        # "conv_w" | "other"


@dataclass
class Bucket:
    index: int
    start: int       # float offset into the flat buffers (inclusive)
    end: int         # exclusive; padded to ALIGN
    names: List[str]

    @property
    def numel(self) -> int:
        return self.end - self.start

    @property
    def nbytes(self) -> int:
        return 4 * self.numel


class FlatLayout:
    """Offsets of every parameter in gradient-ready order plus the bucket plans."""

    def __init__(self, named_params: Sequence[Tuple[str, torch.Tensor]], pad_to: int = 1024):
        self.entries: List[FlatEntry] = []
        off = 0
        for name, p in reversed(list(named_params)):
            kind = "conv_w" if p.dim() == 4 else "other"
            self.entries.append(FlatEntry(name, off, p.numel(), tuple(p.shape), kind))
            off += (p.numel() + ALIGN - 1) // ALIGN * ALIGN
        self.used = off
        self.total = (off + pad_to - 1) // pad_to * pad_to
        self.by_name: Dict[str, FlatEntry] = {e.name: e for e in self.entries}

    def buckets(self, plan: str = "layer") -> List[Bucket]:
        """``layer``: one bucket per conv layer in gradient-ready order, the small tensors
        folded into the neighbouring conv's bucket.  ``ddp``: torch DDP's size
        caps (1 MiB first bucket, 25 MiB afterwards; ref: master/part3/part3.py:116 and the
        probe in SURVEY.md 2.9).  ``single``: everything in one bucket (DDP's first iteration).
        ``cap:<MiB>``: greedy buckets of at least that size, closed at conv-layer boundaries.
        Suffix ``+last`` (any plan): the first conv block -- the last gradients of backward -- gets a bucket of
        its own, so that the only collective that cannot overlap backward moves a few KB."""
        split_last = plan.endswith("+last")
        if split_last:
            plan = plan[: -len("+last")]
        groups: List[List[FlatEntry]] = []
        if plan == "single":
            groups = [list(self.entries)]
        elif plan == "layer":
            cur: List[FlatEntry] = []
            for e in self.entries:
                cur.append(e)
                if e.kind == "conv_w":
                    groups.append(cur)
                    cur = []
            if cur:
                groups[-1].extend(cur)
        elif plan == "ddp":
            cap, cur, size = 1 << 20, [], 0
            for e in self.entries:
                cur.append(e)
                size += 4 * e.numel
                if size >= cap:
                    groups.append(cur)
                    cur, size, cap = [], 0, 25 << 20
            if cur:
                groups.append(cur)
        elif plan.startswith("cap:"):
            # greedy grouping in gradient-ready order, a bucket closes at the first conv weight that takes
            # it to <n> MiB or more (e.g. cap:12 -> 18.9 / 14.2 / 3.9 MB on VGG-11): fewer launches and
            # barrier round trips than "layer", earlier readiness than "ddp"
            try:
                cap = float(plan[4:]) * (1 << 20)
            except ValueError:
                raise ValueError(f"bucket plan {plan!r}: expected cap:<MiB>") from None
            if cap <= 0:
                raise ValueError(f"bucket plan {plan!r}: the cap must be positive")
            cur, size = [], 0
            for e in self.entries:
                cur.append(e)
                size += 4 * e.numel
                if e.kind == "conv_w" and size >= cap:
                    groups.append(cur)
                    cur, size = [], 0
            if cur:
                if groups and not any(e.kind == "conv_w" for e in cur):
                    groups[-1].extend(cur)
                else:
                    groups.append(cur)
        else:
            raise ValueError(f"unknown bucket plan {plan!r}")
        if split_last:
            g = groups[-1]
            convs = [j for j, e in enumerate(g) if e.kind == "conv_w"]
            if len(convs) >= 2:
                cut = convs[-2] + 1          # everything after the second-to-last conv weight of the group
                groups[-1:] = [g[:cut], g[cut:]]
        out: List[Bucket] = []
        for i, g in enumerate(groups):
            start = g[0].offset
            end = self.total if i == len(groups) - 1 else groups[i + 1][0].offset
            out.append(Bucket(i, start, end, [e.name for e in g]))
        return out


class _Block:
    """Workspaces and parameter views of one conv -> BN -> ReLU (-> pool) block."""
    pass


_TUNE_CACHE: Dict[tuple, object] = {}   # (kind, B, hw, cin, cout, taps) -> best tile config, per process


def _igemm_candidates(m_tiles: int, n_total: int, total_kb: int):
    """(block_n, split_k, cta_pair) triples worth timing: every N tile, K splits while the grid is
    small, and the tcgen05 cta_group::2 kernel wherever the tile count allows pairing."""
    out = []
    for bn in (32, 64, 128, 256):
        if n_total % bn:
            continue
        tiles = m_tiles * (n_total // bn)
        pairs = (False, True) if (m_tiles % 2 == 0 and bn >= 64) else (False,)
        for pair in pairs:
            out.append((bn, 1, pair))
            for sk in (2, 3, 4, 6, 9):
                if tiles < 148 and tiles * sk <= 640 and total_kb // sk >= 4 and bn >= 64:
                    out.append((bn, sk, pair))
    return out


def _time_captured(fn, reps: int = 10, replays: int = 5) -> float:
    """Microseconds per call of ``fn``, measured by replaying a CUDA graph of ``reps`` launches
    (removes the host launch cost, which would otherwise hide the differences between ~10 us kernels)."""
    fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    best = float("inf")
    for _ in range(replays):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e3 / reps)
    return best


# tests/test_cpu_executor_dry_run.py sets this to walk the executor's control flow on CPU with the kernel
# wrappers and the CUDA stream / event API replaced by recorders
_DRY_RUN = False
_ACT_DTYPE = torch.bfloat16   # activations / operand copies; the CPU emulation test runs the executor in fp32


class NativeVGG:
    def __init__(self, module: nn.Module, max_batch: int, heap=None, conv_stats_fused: bool = True,
                 autotune: bool = True, dtype: str = "bf16"):
        dev = next(module.parameters()).device
        if dev.type != "cuda" and not _DRY_RUN:
            raise RuntimeError("NativeVGG needs the module on a CUDA device")
        if not K._lib.available():
            raise RuntimeError("libb200dp.so is not available; the native sm_100a path cannot run")
        if dtype not in ("bf16", "tf32"):
            raise ValueError(f"dtype must be 'bf16' or 'tf32', got {dtype!r}")
        # bf16: bf16 activations and operand copies, tcgen05 kind::f16.  tf32: fp32 activations in memory, the
        # tensor cores read them as tf32 (kind::tf32) -- the precision of the reference's fp32 model
        # (ref: master/part1/model.py:18-25 under torch's TF32 convolutions).  Master weights, gradients and
        # the optimizer are fp32 either way.
        self.dtype = dtype
        self.tf32 = dtype == "tf32"
        self.act_dtype = torch.float32 if self.tf32 else _ACT_DTYPE
        self.kblock = 32 if self.tf32 else 64      # channels per 128-byte operand row = K extent of a pipeline stage
        self.module = module
        self.device = dev
        self.max_batch = max_batch
        self.heap = heap
        self.conv_stats_fused = conv_stats_fused
        self.wgrad_push = None        # {block index: conv_wgrad_push arguments}, set by BucketedOverlapSync
        self.defer_head = False       # set by the engine when every training forward is followed by loss_fwd_bwd
        self._head_pending = False
        self._head_done = False
        self.specs = module.block_specs()
        self.layout = FlatLayout(list(module.named_parameters()))
        alloc = (lambda n, dt=torch.float32: heap.zeros(n, dt)) if heap is not None else \
                (lambda n, dt=torch.float32: torch.zeros(n, device=dev, dtype=dt))
        self._alloc = alloc
        n = self.layout.total
        self.flat_params = alloc(n)
        self.flat_grads = alloc(n)
        self.flat_momentum = alloc(n)
        # BatchNorm buffers, flat so that one kernel can broadcast them (part3 semantics)
        self.bn_channels = [s.cout for s in self.specs]
        # running_mean / running_var of every layer followed by the int64 num_batches_tracked counters, in
        # ONE region so that a single kernel broadcasts all BatchNorm state
        n_bn = 2 * sum(self.bn_channels)
        n_nbt = (len(self.specs) + 1) // 2 * 2              # keeps the region a multiple of 16 bytes
        self.bn_state = alloc(n_bn + 2 * n_nbt)             # fp32 words; the tail is viewed as int64
        self.flat_bn = self.bn_state[:n_bn]
        self.flat_nbt_padded = self.bn_state[n_bn:].view(torch.int64)
        self.flat_nbt = self.flat_nbt_padded[: len(self.specs)]
        self.on_bucket_ready: Optional[Callable[[int], None]] = None
        self._bucket_of_block: Dict[int, List[int]] = {}
        self._wgrad_stream = torch.cuda.Stream(dev)
        self._wgrad_fork = [torch.cuda.Event() for _ in self.specs]
        self._wgrad_join = torch.cuda.Event()
        self._dgrad_done = [torch.cuda.Event() for _ in self.specs]
        self._block_prep = None
        self._adopt_parameters()
        self._alloc_workspaces()
        self.refresh_shadows()
        if autotune:
            self.autotune(max_batch)

    # ------------------------------------------------------------------ parameters
    def _view(self, flat: torch.Tensor, e: FlatEntry) -> torch.Tensor:
        t = flat[e.offset: e.offset + e.numel]
        if e.kind == "conv_w":
            co, ci, kh, kw = e.shape
            return t.view(co, kh, kw, ci).permute(0, 3, 1, 2)  # OIHW shape, KRSC memory
        return t.view(e.shape)

    def _adopt_parameters(self) -> None:
        """Copy the module's values into the flat buffers and re-point .data/.grad at views."""
        params = dict(self.module.named_parameters())
        with torch.no_grad():
            for e in self.layout.entries:
                p = params[e.name]
                v = self._view(self.flat_params, e)
                v.copy_(p.data)
                p.data = v
                p.grad = self._view(self.flat_grads, e)
            off = 0
            for i, s in enumerate(self.specs):
                bn = self.module.layers[s.bn_idx]
                rm = self.flat_bn[off: off + s.cout]
                rv = self.flat_bn[off + s.cout: off + 2 * s.cout]
                rm.copy_(bn.running_mean)
                rv.copy_(bn.running_var)
                bn.running_mean.data = rm
                bn.running_var.data = rv
                self.flat_nbt[i].copy_(bn.num_batches_tracked)
                bn.num_batches_tracked.data = self.flat_nbt[i]
                off += 2 * s.cout

    def refresh_from_module(self) -> None:
        """Call after parameters were changed from outside (load_state_dict)."""
        self.refresh_shadows()

    def grad_views_intact(self) -> bool:
        params = dict(self.module.named_parameters())
        return all(params[e.name].grad is not None and
                   params[e.name].grad.data_ptr() == self.flat_grads.data_ptr() + 4 * e.offset
                   for e in self.layout.entries)

    def restore_grad_views(self) -> None:
        params = dict(self.module.named_parameters())
        for e in self.layout.entries:
            params[e.name].grad = self._view(self.flat_grads, e)

    # ------------------------------------------------------------------ workspaces
    def _alloc_workspaces(self) -> None:
        B, dev = self.max_batch, self.device
        bf = self.act_dtype
        L = self.layout
        self.blocks: List[_Block] = []
        # fp32 partial-tile workspace of the split-K convs (kept zero by their finalize kernel)
        need = [B * s.hw * s.hw * max(s.cout, s.cin if i else 0) for i, s in enumerate(self.specs)
                if (B * s.hw * s.hw + 127) // 128 < 148]      # only grids below one wave are ever split
        self.splitk_ws = torch.zeros(max(need + [8]), device=dev)
        zero_floats = sum(4 * s.cout for s in self.specs)
        self.zero_region = torch.zeros(zero_floats, device=dev)  # conv stats + bn-bwd sums, cleared per step
        zoff = 0
        for i, s in enumerate(self.specs):
            b = _Block()
            b.spec = s
            b.first = (i == 0)
            b.fwd_block_n = b.dgrad_block_n = b.wgrad_split_k = 0   # 0 = library heuristic; see autotune()
            b.fwd_split_k = b.dgrad_split_k = 1
            b.wgrad_block_n = 0
            b.fwd_pair = b.dgrad_pair = False
            b.kin = self.kblock if b.first else s.cin   # GEMM K per tap (layer 0: 27 padded to one operand row, 1 tap)
            b.taps = 1 if b.first else 9
            hw = s.hw
            if b.first:
                b.x_in = torch.zeros(B, hw, hw, b.kin, device=dev, dtype=bf)
            else:
                b.x_in = self.blocks[-1].out
            b.y = torch.empty(B, hw, hw, s.cout, device=dev, dtype=bf)
            ho = hw // 2 if s.pool else hw
            b.out = torch.empty(B, ho, ho, s.cout, device=dev, dtype=bf)
            b.dy = torch.empty_like(b.y)
            b.dout = torch.empty_like(b.out)            # gradient w.r.t. the block output
            b.stats = self.zero_region[zoff: zoff + 2 * s.cout]
            b.sums = self.zero_region[zoff + 2 * s.cout: zoff + 4 * s.cout]
            zoff += 4 * s.cout
            b.mean = torch.empty(s.cout, device=dev)
            b.invstd = torch.empty(s.cout, device=dev)
            kpad = b.taps * b.kin
            b.wd = None if b.first else torch.empty(s.cin, 9 * s.cout, device=dev, dtype=bf)
            b._wf_unpadded = self.tf32 and not b.first     # fp32 master [Cout][9*Cin] IS the forward operand
            b.wf = None if b._wf_unpadded else torch.empty(s.cout, kpad, device=dev, dtype=bf)
            ew, eb = L.by_name[f"layers.{s.conv_idx}.weight"], L.by_name[f"layers.{s.conv_idx}.bias"]
            eg, ebt = L.by_name[f"layers.{s.bn_idx}.weight"], L.by_name[f"layers.{s.bn_idx}.bias"]
            fp, fg = self.flat_params, self.flat_grads
            sl = lambda flat, e: flat[e.offset: e.offset + e.numel]
            b.w_master, b.bias, b.gamma, b.beta = sl(fp, ew), sl(fp, eb), sl(fp, eg), sl(fp, ebt)
            # what the weight re-layout kernels write for this block: (master, Wf or None, Wd or None, ...)
            b.prep = (b.w_master, b.wf, b.wd, s.cout, s.cin, 9, kpad)
            if b._wf_unpadded:
                b.wf = b.w_master.view(s.cout, kpad)
            b.g_w, b.g_bias, b.g_gamma, b.g_beta = sl(fg, ew), sl(fg, eb), sl(fg, eg), sl(fg, ebt)
            bn = self.module.layers[s.bn_idx]
            b.running_mean, b.running_var, b.nbt = bn.running_mean.data, bn.running_var.data, self.flat_nbt[i:i + 1]
            self.blocks.append(b)
        efw, efb = L.by_name["fc1.weight"], L.by_name["fc1.bias"]
        self.fc_w = self.flat_params[efw.offset: efw.offset + efw.numel].view(efw.shape)
        self.fc_b = self.flat_params[efb.offset: efb.offset + efb.numel]
        self.g_fc_w = self.flat_grads[efw.offset: efw.offset + efw.numel].view(efw.shape)
        self.g_fc_b = self.flat_grads[efb.offset: efb.offset + efb.numel]
        nc = efw.shape[0]
        self.logits = torch.zeros(B, nc, device=dev)
        self.dlogits = torch.zeros(B, nc, device=dev)
        self.loss = torch.zeros((), device=dev)
        self.correct = torch.zeros((), device=dev, dtype=torch.int32)

    def set_bucket_plan(self, buckets: Sequence[Bucket]) -> None:
        """Tell the executor which bucket becomes complete after which block's weight gradient."""
        self._bucket_of_block = {}
        L = self.layout

        def block_of(name: str) -> int:
            if name.startswith("layers."):
                idx = int(name.split(".")[1])
                for i, b in enumerate(self.blocks):
                    if idx in (b.spec.conv_idx, b.spec.bn_idx):
                        return i
            return len(self.blocks) - 1  # fc1.* gradients exist before the last block's backward

        for bk in buckets:
            # complete once the lowest layer that owns a tensor of the bucket has run its backward
            self._bucket_of_block.setdefault(min(block_of(n) for n in bk.names), []).append(bk.index)
        self.buckets = list(buckets)

    # ------------------------------------------------------------------ tile autotuning
    def autotune(self, B: int) -> None:
        """Pick the N tile of every forward / dgrad conv and the split-K of every wgrad by timing the
        candidates on this GPU at batch ``B`` (the shapes are tiny and L2->SM bandwidth bound, so the
        best tile depends on how many CTAs a shape yields; a static rule leaves 5-10% on the table)."""
        side = torch.cuda.Stream(self.device)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for i, b in enumerate(self.blocks):
                s = b.spec
                x_in, y, dy = b.x_in[:B], b.y[:B], b.dy[:B]
                key = (self.dtype, B, s.hw, b.kin, s.cout, b.taps)
                m_tiles = (B * s.hw * s.hw + 127) // 128
                if ("fwd",) + key not in _TUNE_CACHE:
                    t = {c: _time_captured(lambda c=c: K.conv_igemm(x_in, b.wf, b.bias, y, stats=b.stats, taps=b.taps,
                                                                    block_n=c[0], split_k=c[1], cta_pair=c[2],
                                                                    workspace=self.splitk_ws))
                         for c in _igemm_candidates(m_tiles, s.cout, b.taps * b.kin // self.kblock)}
                    _TUNE_CACHE[("fwd",) + key] = min(t, key=t.get)
                b.fwd_block_n, b.fwd_split_k, b.fwd_pair = _TUNE_CACHE[("fwd",) + key]
                if ("wgrad",) + key not in _TUNE_CACHE:
                    tiles_n = [0] + ([256] if b.kin % 256 == 0 else [])
                    t = {(k, n): _time_captured(lambda k=k, n=n: K.conv_wgrad(dy, x_in, b.g_w, taps=b.taps, split_k=k,
                                                                              cin_valid=(27 if b.first else None),
                                                                              block_n=n))
                         for k in (0, 1, 2, 4, 8, 16) for n in tiles_n}
                    _TUNE_CACHE[("wgrad",) + key] = min(t, key=t.get)
                b.wgrad_split_k, b.wgrad_block_n = _TUNE_CACHE[("wgrad",) + key]
                if i > 0:
                    dprev = self.blocks[i - 1].dout[:B]
                    if ("dgrad",) + key not in _TUNE_CACHE:
                        t = {c: _time_captured(lambda c=c: K.conv_igemm(dy, b.wd, None, dprev, block_n=c[0],
                                                                        split_k=c[1], cta_pair=c[2],
                                                                        workspace=self.splitk_ws))
                             for c in _igemm_candidates(m_tiles, s.cin, 9 * s.cout // self.kblock)}
                        _TUNE_CACHE[("dgrad",) + key] = min(t, key=t.get)
                    b.dgrad_block_n, b.dgrad_split_k, b.dgrad_pair = _TUNE_CACHE[("dgrad",) + key]
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.flat_grads.zero_()
        self.zero_region.zero_()

    # ------------------------------------------------------------------ weight shadows
    def refresh_shadows(self) -> None:
        """bf16 operand copies of the fp32 master weights: Wf (forward) and Wd (dgrad)."""
        if getattr(self, "_prep", None) is None:
            self._prep = K.BatchedWeightPrep([b.prep for b in self.blocks])
        self._prep()

    def prep_block(self, i: int) -> None:
        """Refresh the bf16 weight copies of block ``i`` on the current stream, after the data gradient
        of that block (the last reader of the old copies in this step) has finished."""
        if self._block_prep is None:
            self._block_prep = [K.BatchedWeightPrep([b.prep]) for b in self.blocks]
        torch.cuda.current_stream().wait_event(self._dgrad_done[i])
        self._block_prep[i]()

    def blocks_of_bucket(self, k: int) -> List[int]:
        return [i for i, ks in self._bucket_of_block.items() if k in ks]

    @staticmethod
    def _pair_ok(pair: bool, B: int, hw: int) -> bool:
        """The CTA-pair kernels need an even number of 128-pixel tiles: a ragged last batch (the tile shapes
        were tuned at the full batch) falls back to the one-CTA kernel with the same N tile."""
        return bool(pair) and ((B * hw * hw + 127) // 128) % 2 == 0

    # ------------------------------------------------------------------ forward
    def forward(self, x: torch.Tensor, training: bool = True) -> torch.Tensor:
        """x: fp32 NCHW [B,3,32,32] on the device.  Returns the logits workspace view [B,10]."""
        B = x.shape[0]
        if B > self.max_batch:
            raise ValueError(f"batch {B} exceeds the executor's max_batch {self.max_batch}")
        self._B = B
        if training:
            K.fill_zero(self.zero_region)
        for b in self.blocks:
            s = b.spec
            y = b.y[:B]
            fused = training and self.conv_stats_fused
            x_in = b.x_in[:B]
            if b.first:
                K.im2col3x3_c3(x.contiguous(), x_in, upper_half_zeroed=True)   # x_in is allocated as zeros
            K.conv_igemm(x_in, b.wf, b.bias, y, stats=b.stats if fused else None, taps=b.taps,
                         block_n=b.fwd_block_n, split_k=b.fwd_split_k, cta_pair=self._pair_ok(b.fwd_pair, B, s.hw),
                         workspace=self.splitk_ws)
            if training and not fused:
                K.bn_stats(y, b.stats)
            K.bn_relu_pool_fwd(y, b.stats, b.gamma, b.beta, b.running_mean, b.running_var, b.nbt, b.mean, b.invstd,
                               b.out[:B], s.pool, training)
        feat = self.blocks[-1].out[:B]
        if training and self.defer_head:
            self._head_pending = True     # the caller promises loss_fwd_bwd(): the fused head fills the logits
        else:
            self._head_pending = False
            K.linear_fwd(feat, self.fc_w, self.fc_b, self.logits[:B])
        return self.logits[:B]

    def loss_fwd_bwd(self, target: torch.Tensor, grad_scale: float = 1.0, need_grad: bool = True) -> torch.Tensor:
        """Mean cross-entropy of the last forward's logits; fills dlogits = dL/dlogits * grad_scale."""
        B = self._B
        if self._head_pending:
            self._head_pending = False
            last = self.blocks[-1]
            if need_grad:
                self._clear_grads()
                K.head_fwd_bwd(last.out[:B], self.fc_w, self.fc_b, target, self.logits[:B], self.loss, self.correct,
                               self.dlogits[:B], last.dout[:B], self.g_fc_w, self.g_fc_b, grad_scale)
                self._head_done = True
                return self.loss
            K.linear_fwd(last.out[:B], self.fc_w, self.fc_b, self.logits[:B])
        K.cross_entropy(self.logits[:B], target, self.loss, self.dlogits[:B] if need_grad else None, self.correct,
                        grad_scale)
        return self.loss

    def _clear_grads(self) -> None:
        # split-K weight gradients are accumulated with red.global.add; the conv biases rely on this zero too
        K.fill_zero(self.flat_grads)

    # ------------------------------------------------------------------ backward
    def backward(self, dlogits: Optional[torch.Tensor] = None) -> None:
        """Back-propagate from dlogits (default: the buffer filled by ``loss_fwd_bwd``) and write
        every parameter gradient into ``flat_grads``.  Calls ``on_bucket_ready(k)`` as soon as the
        last gradient of bucket k has been *launched* on the current stream."""
        B = self._B
        dl = self.dlogits[:B] if dlogits is None else dlogits
        last = self.blocks[-1]
        main = torch.cuda.current_stream()
        side = self._wgrad_stream
        if self._head_done and dlogits is None:
            self._head_done = False     # the fused head already cleared the gradients and ran the fc backward
        else:
            self._head_done = False
            self._clear_grads()
            K.linear_bwd(last.out[:B], self.fc_w, dl, self.g_fc_w, self.g_fc_b, last.dout[:B])
        reduced = False   # True: the dgrad of the block above already accumulated this block's sums
        for i in range(len(self.blocks) - 1, -1, -1):
            b = self.blocks[i]
            s = b.spec
            K.bn_relu_pool_bwd(b.y[:B], b.dout[:B], b.mean, b.invstd, b.gamma, b.beta, b.sums, b.dy[:B], b.g_gamma,
                               b.g_beta, s.pool, out=b.out[:B], skip_reduce=reduced)
            reduced = False
            # the conv bias feeds a training-mode BatchNorm, so its exact gradient is zero:
            # it simply keeps the zero written by the flat_grads clear at the top of backward().
            #
            # wgrad(i) and dgrad(i) both only read dy(i) and are independent: the weight gradient goes to
            # a second stream so that it runs next to the data-gradient chain (each of these GEMMs keeps
            # the tensor pipe ~40% busy on its own).  The bucket-ready signal is recorded on that stream.
            self._wgrad_fork[i].record(main)
            if i > 0:
                # the dgrad epilogue also accumulates the BatchNorm-backward sums of the block below,
                # which then skips its own reduction pass
                prev = self.blocks[i - 1]
                K.conv_dgrad_bn_reduce(b.dy[:B], b.wd, prev.dout[:B], prev.out[:B], prev.gamma, prev.beta, prev.sums,
                                       block_n=b.dgrad_block_n, split_k=b.dgrad_split_k,
                                       cta_pair=self._pair_ok(b.dgrad_pair, B, s.hw), workspace=self.splitk_ws)
                reduced = True
            self._dgrad_done[i].record(main)   # after this the layer's bf16 weight copies may be rewritten
            side.wait_event(self._wgrad_fork[i])
            with torch.cuda.stream(side):
                if self.wgrad_push is not None and i in self.wgrad_push:
                    # optional path: GEMM + transfer half of the reduce-scatter in one kernel
                    K.conv_wgrad_push(b.dy[:B], b.x_in[:B], b.g_w, split_k=b.wgrad_split_k, block_n=b.wgrad_block_n,
                                      **self.wgrad_push[i])
                else:
                    K.conv_wgrad(b.dy[:B], b.x_in[:B], b.g_w, taps=b.taps, cin_valid=(27 if b.first else None),
                                 split_k=b.wgrad_split_k, block_n=b.wgrad_block_n)
                if self.on_bucket_ready is not None:
                    for k in self._bucket_of_block.get(i, ()):
                        self.on_bucket_ready(k)
        self._wgrad_join.record(side)
        main.wait_event(self._wgrad_join)

    # ------------------------------------------------------------------ torch.autograd bridge
    def autograd_forward(self, x: torch.Tensor, training: bool) -> torch.Tensor:
        params = [p for _, p in self.module.named_parameters()]
        return _NativeVGGFunction.apply(x, self, training, *params)


class _NativeVGGFunction(torch.autograd.Function):
    """Lets ``loss.backward()`` drive the native backward.  Gradients are *written* into the
    flat gradient buffer (the tutorial zeroes gradients every iteration, so overwrite ==
    accumulate-after-zero; ref: master/part1/part1.py:34)."""

    @staticmethod
    def forward(ctx, x, native: NativeVGG, training: bool, *params):
        ctx.native = native
        ctx.training = training
        logits = native.forward(x, training)
        return logits.clone()

    @staticmethod
    def backward(ctx, dlogits):
        native: NativeVGG = ctx.native
        if not ctx.training:
            raise RuntimeError("backward through an eval-mode native forward is not supported")
        native.backward(dlogits.contiguous().float())
        grads = []
        params = dict(native.module.named_parameters())
        for name, p in native.module.named_parameters():
            e = native.layout.by_name[name]
            view = native._view(native.flat_grads, e)
            if p.grad is not None and p.grad.data_ptr() == view.data_ptr():
                grads.append(None)      # already in place; returning it would double it
            else:
                grads.append(view)      # autograd installs the view as p.grad
        return (None, None, None, *grads)
