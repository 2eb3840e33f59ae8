"""Thin torch-tensor wrappers over the native sm_100a kernels.

Each function validates dtypes/layouts, allocates nothing unless documented and
launches on the current CUDA stream, so the whole step can be captured in a CUDA
graph.  Activations are bf16 NHWC ``[B, H, W, C]``; see ``include/b200dp.h``.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import _lib
from ._lib import call, ptr, stream_ptr

BN_EPS = 1e-5
BN_MOMENTUM = 0.1


def _chk(t: torch.Tensor, dtype: torch.dtype, name: str) -> None:
    if not t.is_cuda or t.dtype != dtype or not t.is_contiguous():
        raise ValueError(f"{name}: expected contiguous CUDA {dtype} tensor, got {t.dtype} on {t.device} "
                         f"contiguous={t.is_contiguous()}")


def _act_dtype(t: torch.Tensor, name: str) -> int:
    """0 for bf16 activations, 1 for fp32 activations (tf32 tensor-core math); checks layout."""
    if t.dtype not in (torch.bfloat16, torch.float32):
        raise ValueError(f"{name}: activations are bf16 or fp32 (tf32 path), got {t.dtype}")
    _chk(t, t.dtype, name)
    return 1 if t.dtype == torch.float32 else 0


def conv_igemm(act: torch.Tensor, wgt: torch.Tensor, bias: Optional[torch.Tensor], out: torch.Tensor,
               stats: Optional[torch.Tensor] = None, taps: int = 9, block_n: int = 0, split_k: int = 1,
               workspace: Optional[torch.Tensor] = None, cta_pair: bool = False) -> torch.Tensor:
    """out[B,H,W,Cout] = conv3x3(act[B,H,W,Cin], wgt[Cout, taps*Cin]) + bias (tcgen05 implicit GEMM).

    ref: master/part1/model.py:18-23.  Also the dgrad kernel when fed dY and the flipped/transposed weights.
    bf16 tensors run ``tcgen05.mma kind::f16``; fp32 tensors (``act``, ``wgt`` and ``out`` alike) run
    ``kind::tf32`` on the fp32 data -- the reference's precision."""
    dt = _act_dtype(act, "act")
    _chk(wgt, act.dtype, "wgt")
    _chk(out, act.dtype, "out")
    B, H, W, Cin = act.shape
    Cout = wgt.shape[0]
    assert wgt.shape[1] == taps * Cin, (wgt.shape, taps, Cin)
    assert out.shape == (B, H, W, Cout)
    if split_k > 1:
        # K loop split over several CTAs per tile; `workspace` (fp32, >= B*H*W*Cout, zero) holds the partials
        assert workspace is not None and workspace.dtype == torch.float32 and workspace.numel() >= out.numel()
    call("b200dp_conv_igemm_ex", ptr(act), ptr(wgt), ptr(bias), ptr(out), ptr(stats),
         ptr(workspace) if split_k > 1 else None, B, H, W, Cin, Cout, taps, block_n, split_k, int(bool(cta_pair)), dt,
         stream_ptr())
    return out


def conv_dgrad_bn_reduce(dy: torch.Tensor, wd: torch.Tensor, dout_prev: torch.Tensor, out_prev: torch.Tensor,
                         gamma: torch.Tensor, beta: torch.Tensor, sums: torch.Tensor, block_n: int = 0,
                         split_k: int = 1, workspace: Optional[torch.Tensor] = None,
                         cta_pair: bool = False) -> torch.Tensor:
    """Data gradient of a 3x3 conv fused with the BatchNorm-backward reduction of the block below:
    ``dout_prev = dgrad(dy, wd)`` and ``sums += [sum_{a>0} g, sum_{a>0} g*(a-beta)/gamma]`` where
    ``a = out_prev`` is that block's stored output.  Saves the ``bn_bwd_reduce_from_out`` launch and its
    read of ``dout_prev`` (ref: the autograd of master/part1/model.py:18-23 run by part1.py:37)."""
    dt = _act_dtype(dy, "dy")
    for t, n in ((wd, "wd"), (dout_prev, "dout_prev"), (out_prev, "out_prev")):
        _chk(t, dy.dtype, n)
    B, H, W, Cin = dy.shape
    Cout = wd.shape[0]
    assert wd.shape[1] == 9 * Cin and dout_prev.shape == (B, H, W, Cout) and out_prev.shape == dout_prev.shape
    assert sums.numel() >= 2 * Cout and gamma.numel() == Cout and beta.numel() == Cout
    if split_k > 1:
        assert workspace is not None and workspace.dtype == torch.float32 and workspace.numel() >= dout_prev.numel()
    call("b200dp_conv_dgrad_bn_reduce_ex", ptr(dy), ptr(wd), ptr(dout_prev), ptr(workspace) if split_k > 1 else None,
         ptr(out_prev), ptr(gamma), ptr(beta), ptr(sums), B, H, W, Cin, Cout, 9, block_n, split_k, int(bool(cta_pair)),
         dt, stream_ptr())
    return dout_prev


def conv_wgrad(dy: torch.Tensor, x: torch.Tensor, dw: torch.Tensor, taps: int = 9, cin_valid: Optional[int] = None,
               split_k: int = 0, accumulate: bool = False, block_n: int = 0) -> torch.Tensor:
    """dw[Cout, taps, cin_valid] (fp32) (+)= sum_p dy[p, Cout] * x[p+tap, Cin].  With split-K or
    ``accumulate`` the partial tiles are added onto ``dw`` (zero it first for a plain gradient).
    bf16 or fp32 (tf32 math) operands, as for ``conv_igemm``."""
    dt = _act_dtype(dy, "dy")
    _chk(x, dy.dtype, "x")
    _chk(dw, torch.float32, "dw")
    B, H, W, Cout = dy.shape
    Cin = x.shape[3]
    cv = Cin if cin_valid is None else cin_valid
    assert dw.numel() == Cout * taps * cv
    call("b200dp_conv_wgrad_ex", ptr(dy), ptr(x), ptr(dw), B, H, W, Cin, Cout, taps, cv, split_k, int(accumulate),
         block_n, dt, stream_ptr())
    return dw


def conv_wgrad_push(dy: torch.Tensor, x: torch.Tensor, dw: torch.Tensor, *, counters: torch.Tensor, world: int,
                    rank: int, peer_stage, flat_off: int, own_start: int, own_len: int, stage_base: int,
                    split_k: int = 0, block_n: int = 0) -> torch.Tensor:
    """Weight gradient whose epilogue also forwards every finished tile to the rank that owns it in the
    data-parallel reduce-scatter (``peer_stage``: ctypes array of ``world`` staging-buffer pointers).
    ``dw`` must be zero on entry and sit at element ``flat_off`` of the flat gradient buffer."""
    _chk(dy, torch.bfloat16, "dy")
    _chk(x, torch.bfloat16, "x")
    _chk(dw, torch.float32, "dw")
    assert counters.dtype == torch.int32 and counters.is_cuda
    B, H, W, Cout = dy.shape
    Cin = x.shape[3]
    assert dw.numel() == Cout * 9 * Cin
    call("b200dp_conv_wgrad_push", ptr(dy), ptr(x), ptr(dw), B, H, W, Cin, Cout, 9, split_k, block_n, ptr(counters),
         counters.numel(), world, rank, peer_stage, flat_off, own_start, own_len, stage_base, stream_ptr())
    return dw


def im2col3x3_c3(x_nchw: torch.Tensor, a0: torch.Tensor, upper_half_zeroed: bool = False) -> torch.Tensor:
    """Layer-0 patch gather: bf16 rows of 64 columns or fp32 rows of 32 columns (27 real ones either way).
    ``upper_half_zeroed``: columns 32..63 of the bf16 matrix already hold their zeros and are not rewritten."""
    _chk(x_nchw, torch.float32, "x")
    dt = _act_dtype(a0, "a0")
    B, C, H, W = x_nchw.shape
    assert C == 3 and a0.numel() == B * H * W * (32 if dt else 64)
    call("b200dp_im2col3x3_c3_ex2", ptr(x_nchw), ptr(a0), B, H, W, dt, int(upper_half_zeroed), stream_ptr())
    return a0


def bn_stats(y: torch.Tensor, stats: torch.Tensor) -> torch.Tensor:
    dt = _act_dtype(y, "y")
    _chk(stats, torch.float32, "stats")
    C = y.shape[-1]
    call("b200dp_bn_stats_ex", ptr(y), ptr(stats), y.numel() // C, C, dt, stream_ptr())
    return stats


def bn_relu_pool_fwd(y, stats, gamma, beta, running_mean, running_var, num_batches_tracked, save_mean, save_invstd,
                     out, pool: bool, training: bool, momentum: float = BN_MOMENTUM, eps: float = BN_EPS):
    B, H, W, C = y.shape
    dt = _act_dtype(y, "y")
    _chk(out, y.dtype, "out")
    call("b200dp_bn_relu_pool_fwd_ex", ptr(y), ptr(stats), ptr(gamma), ptr(beta), ptr(running_mean),
         ptr(running_var), ptr(num_batches_tracked), ptr(save_mean), ptr(save_invstd), ptr(out), B, H, W, C, int(pool),
         int(training), momentum, eps, dt, stream_ptr())
    return out


def bn_relu_pool_bwd(y, dout, mean, invstd, gamma, beta, sums, dy, dgamma, dbeta, pool: bool,
                     out: Optional[torch.Tensor] = None, skip_reduce: bool = False):
    """Per-channel reduction into ``sums`` (must be zero), then the dY pass.  With ``out`` (the block's stored
    output) the reduction reads it instead of the full-resolution conv output: 2.5x fewer bytes on pooled
    layers and no arg-max search.  ``skip_reduce``: ``sums`` already holds the reduction
    (``conv_dgrad_bn_reduce`` of the block above produced it), only apply."""
    B, H, W, C = y.shape
    dt = _act_dtype(y, "y")
    _chk(dout, y.dtype, "dout")
    _chk(dy, y.dtype, "dy")
    if not skip_reduce:
        if out is not None:
            call("b200dp_bn_relu_pool_bwd_reduce_from_out_ex", ptr(out), ptr(dout), ptr(gamma), ptr(beta), ptr(sums),
                 dout.numel() // C, C, dt, stream_ptr())
        else:
            call("b200dp_bn_relu_pool_bwd_reduce_ex", ptr(y), ptr(dout), ptr(mean), ptr(invstd), ptr(gamma),
                 ptr(beta), ptr(sums), B, H, W, C, int(pool), dt, stream_ptr())
    call("b200dp_bn_relu_pool_bwd_apply_ex", ptr(y), ptr(dout), ptr(mean), ptr(invstd), ptr(gamma), ptr(beta),
         ptr(sums), ptr(dy), ptr(dgamma), ptr(dbeta), B, H, W, C, int(pool), dt, stream_ptr())
    return dy


def linear_fwd(x, w, b, logits):
    B, Fdim = x.shape[0], x.numel() // x.shape[0]
    call("b200dp_linear_fwd_ex", ptr(x), ptr(w), ptr(b), ptr(logits), B, Fdim, w.shape[0], _act_dtype(x, "x"),
         stream_ptr())
    return logits


def cross_entropy(logits, target, loss=None, dlogits=None, correct=None, grad_scale: float = 1.0):
    B, O = logits.shape
    assert target.dtype == torch.int64
    call("b200dp_cross_entropy", ptr(logits), ptr(target), ptr(loss), ptr(dlogits), ptr(correct), B, O,
         float(grad_scale), stream_ptr())


def head_fwd_bwd(x, w, b, target, logits, loss, correct, dlogits, dx, dw, db, grad_scale: float = 1.0):
    """Classifier head of a training step in one launch: ``linear_fwd`` + ``cross_entropy`` + ``linear_bwd``
    (ref: master/part1/model.py:44-46 and part1.py:35-37).  ``dw`` / ``db`` are accumulated onto."""
    B, Fdim = x.shape[0], x.numel() // x.shape[0]
    assert target.dtype == torch.int64 and dx.dtype == x.dtype
    call("b200dp_head_fwd_bwd_ex", ptr(x), ptr(w), ptr(b), ptr(target), ptr(logits), ptr(loss), ptr(correct),
         ptr(dlogits), ptr(dx), ptr(dw), ptr(db), B, Fdim, w.shape[0], float(grad_scale), _act_dtype(x, "x"),
         stream_ptr())


def linear_bwd(x, w, dlogits, dw, db, dx):
    B, Fdim = x.shape[0], x.numel() // x.shape[0]
    assert dx.dtype == x.dtype
    call("b200dp_linear_bwd_ex", ptr(x), ptr(w), ptr(dlogits), ptr(dw), ptr(db), ptr(dx), B, Fdim, w.shape[0],
         _act_dtype(x, "x"), stream_ptr())


def sgd_momentum(p, g, buf, lr: float, momentum: float, wd: float, grad_scale: float = 1.0):
    """Fused SGD on flat fp32 buffers (ref: master/part1/part1.py:38,98-99)."""
    assert p.numel() == g.numel() == buf.numel()
    call("b200dp_sgd_momentum", ptr(p), ptr(g), ptr(buf), p.numel(), lr, momentum, wd, grad_scale, stream_ptr())


def fill_zero(t: torch.Tensor) -> torch.Tensor:
    """Clear a contiguous tensor whose byte size is a multiple of 16 (flat buffers are padded to that)."""
    call("b200dp_fill_zero", ptr(t), t.numel() * t.element_size(), stream_ptr())
    return t


def copy_pair(dst0: torch.Tensor, src0: torch.Tensor, dst1: torch.Tensor, src1: torch.Tensor) -> None:
    """``dst0 <- src0`` and ``dst1 <- src1`` (device to device, contiguous, multiples of 16 bytes) in one launch."""
    n0, n1 = src0.numel() * src0.element_size(), src1.numel() * src1.element_size()
    assert dst0.numel() * dst0.element_size() >= n0 and dst1.numel() * dst1.element_size() >= n1
    call("b200dp_copy_pair", ptr(dst0), ptr(src0), n0, ptr(dst1), ptr(src1), n1, stream_ptr())


def _operand_dtype(wf, wd) -> int:
    t = wf if wf is not None else wd
    return 1 if t.dtype == torch.float32 else 0


def prep_conv_weights(master, wf, wd, Cout: int, Cin: int, taps: int, kpad: int):
    """Operand copies of one conv layer's fp32 master weights: ``wf`` [Cout, kpad] (forward, zero padded) and
    ``wd`` [Cin, taps*Cout] (data gradient: transposed, taps flipped); either may be None.  bf16 tensors for the
    bf16 kernels, fp32 tensors for the tf32 kernels."""
    call("b200dp_prep_conv_weights_ex", ptr(master), ptr(wf), ptr(wd), Cout, Cin, taps, kpad, _operand_dtype(wf, wd),
         stream_ptr())


class BatchedWeightPrep:
    """All conv layers' operand copies (Wf and the tap-flipped transposed Wd) in ONE launch.  The
    pointer tables are built once; layers that are not multiples of 64 (the 3-channel first conv)
    fall back to the single-layer kernel; layers with nothing to write (tf32 forward operand = the master
    tensor, no data gradient) are skipped."""

    def __init__(self, layers):
        """``layers``: iterable of (master, wf_or_None, wd_or_None, Cout, Cin, taps, kpad)."""
        import ctypes
        layers = [l for l in layers if l[1] is not None or l[2] is not None]
        odd = lambda l: bool(l[3] % 64 or l[4] % 64 or l[6] != l[5] * l[4])
        self._single = [l for l in layers if odd(l)]
        batched = [l for l in layers if not odd(l)]
        self._n = len(batched)
        self._keep = batched
        if self._n:
            vp = ctypes.c_void_p * self._n
            ia = ctypes.c_int * self._n
            self._dtype = _operand_dtype(batched[0][1], batched[0][2])
            self._master = vp(*[l[0].data_ptr() for l in batched])
            self._wf = vp(*[(l[1].data_ptr() if l[1] is not None else None) for l in batched])
            self._wd = vp(*[(l[2].data_ptr() if l[2] is not None else None) for l in batched])
            self._cout, self._cin = ia(*[l[3] for l in batched]), ia(*[l[4] for l in batched])
            self._taps, self._kpad = ia(*[l[5] for l in batched]), ia(*[l[6] for l in batched])

    def __call__(self) -> None:
        for m, wf, wd, co, ci, taps, kpad in self._single:
            prep_conv_weights(m, wf, wd, co, ci, taps, kpad)
        if self._n:
            call("b200dp_prep_conv_weights_batched_ex", self._n, self._master, self._wf, self._wd, self._cout,
                 self._cin, self._taps, self._kpad, self._dtype, stream_ptr())


def version() -> str:
    return _lib.version()
