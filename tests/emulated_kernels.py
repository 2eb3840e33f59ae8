"""torch (CPU, fp32) restatements of the kernel wrappers in ops/native.py, with the same signatures and the
same in-place contracts (which buffer is read, which is overwritten, which is accumulated onto).

Used by tests/test_cpu_executor_emulated.py to run the GPU executor's data flow numerically on the host:
if `NativeVGG.backward()` hands a kernel the wrong buffer, or an experimental toggle changes the maths, the
gradients stop matching torch autograd.  The layouts are the native ones: activations NHWC, conv weights
KRSC, forward operand Wf = [Cout][tap*Cin + ci], data-gradient operand Wd = [Cin][flipped tap * Cout + co],
layer 0 as a 27(->64)-column patch matrix.
"""
import types

import torch
import torch.nn.functional as F

BN_MOMENTUM, BN_EPS = 0.1, 1e-5


def _put(dst, src):
    dst.copy_(src.reshape(dst.shape).to(dst.dtype))
    return dst


class EmulatedKernels:
    def __init__(self):
        self.calls = []
        self._lib = types.SimpleNamespace(available=lambda: True)

    def _log(self, name):
        self.calls.append(name)

    # ---- fills / copies ---------------------------------------------------------------------------
    def fill_zero(self, t):
        self._log("fill_zero")
        return t.zero_()

    def copy_pair(self, dst0, src0, dst1, src1):
        self._log("copy_pair")
        dst0.view(-1)[: src0.numel()].copy_(src0.view(-1))
        dst1.view(-1)[: src1.numel()].copy_(src1.view(-1))

    # ---- layer 0 patch gather --------------------------------------------------------------------
    def im2col3x3_c3(self, x_nchw, a0, upper_half_zeroed=False):
        self._log("im2col3x3_c3")
        B, C, H, W = x_nchw.shape
        xp = F.pad(x_nchw.float(), (1, 1, 1, 1))
        cols = torch.zeros(B, H, W, 64)
        for r in range(3):
            for s in range(3):
                for c in range(3):
                    cols[..., (r * 3 + s) * 3 + c] = xp[:, c, r:r + H, s:s + W]
        return _put(a0, cols)

    # ---- implicit-GEMM conv (forward and data gradient) ----------------------------------------------
    def conv_igemm(self, act, wgt, bias, out, stats=None, taps=9, block_n=0, split_k=1, workspace=None, cta_pair=False):
        self._log("conv_igemm")
        B, H, W, Cin = act.shape
        Cout = wgt.shape[0]
        a = act.float()
        w = wgt.float()
        if taps == 1:
            y = a.reshape(-1, Cin) @ w.t()
            y = y.view(B, H, W, Cout)
        else:
            k = w.view(Cout, 3, 3, Cin).permute(0, 3, 1, 2)              # OIHW
            y = F.conv2d(a.permute(0, 3, 1, 2), k, None, padding=1).permute(0, 2, 3, 1)
        if bias is not None:
            y = y + bias.float().view(1, 1, 1, -1)
        _put(out, y)
        if stats is not None:
            stored = out.float().reshape(-1, Cout)                        # statistics of the values as stored
            stats[:Cout] += stored.sum(0)
            stats[Cout:2 * Cout] += (stored * stored).sum(0)
        return out

    def bn_stats(self, y, stats):
        self._log("bn_stats")
        C = y.shape[-1]
        v = y.float().reshape(-1, C)
        stats[:C] += v.sum(0)
        stats[C:2 * C] += (v * v).sum(0)

    # ---- BatchNorm + ReLU (+ 2x2 max-pool) ---------------------------------------------------------
    def bn_relu_pool_fwd(self, y, stats, gamma, beta, running_mean, running_var, num_batches_tracked, save_mean,
                         save_invstd, out, pool, training, momentum=BN_MOMENTUM, eps=BN_EPS):
        self._log("bn_relu_pool_fwd")
        B, H, W, C = y.shape
        P = B * H * W
        if training:
            mean = stats[:C] / P
            var = (stats[C:2 * C] / P - mean * mean).clamp_min(0)
            invstd = torch.rsqrt(var + eps)
            save_mean.copy_(mean)
            save_invstd.copy_(invstd)
            unbias = P / (P - 1) if P > 1 else 1.0
            running_mean.mul_(1 - momentum).add_(momentum * mean)
            running_var.mul_(1 - momentum).add_(momentum * var * unbias)
            if num_batches_tracked is not None:
                num_batches_tracked += 1
        else:
            mean, invstd = running_mean, torch.rsqrt(running_var + eps)
        z = (y.float() - mean) * invstd * gamma + beta
        a = torch.relu(z)
        if pool:
            a = F.max_pool2d(a.permute(0, 3, 1, 2), 2).permute(0, 2, 3, 1)
        return _put(out, a)

    def bn_relu_pool_bwd(self, y, dout, mean, invstd, gamma, beta, sums, dy, dgamma, dbeta, pool, out=None,
                         skip_reduce=False):
        self._log("bn_relu_pool_bwd")
        B, H, W, C = y.shape
        P = B * H * W
        xhat = (y.float() - mean) * invstd
        z = xhat * gamma + beta
        g = dout.float()
        if pool:
            with torch.enable_grad():          # (the executor may be running inside an autograd backward)
                zp = z.permute(0, 3, 1, 2).detach().requires_grad_(True)
                pooled = F.max_pool2d(torch.relu(zp), 2)
                pooled.backward(g.permute(0, 3, 1, 2))
            dz = zp.grad.permute(0, 2, 3, 1)
        else:
            dz = g * (z > 0)
        if not skip_reduce:
            sums[:C] += dz.reshape(-1, C).sum(0)
            sums[C:2 * C] += (dz * xhat).reshape(-1, C).sum(0)
        s1, s2 = sums[:C], sums[C:2 * C]
        _put(dy, gamma * invstd * (dz - s1 / P - xhat * s2 / P))
        dgamma.copy_(s2)
        dbeta.copy_(s1)
        return dy

    # ---- classifier and loss ------------------------------------------------------------------------
    def linear_fwd(self, x, w, b, logits):
        self._log("linear_fwd")
        return _put(logits, x.float().reshape(x.shape[0], -1) @ w.t() + b)

    def cross_entropy(self, logits, target, loss=None, dlogits=None, correct=None, grad_scale=1.0):
        self._log("cross_entropy")
        lg = logits.float()
        if loss is not None:
            loss.copy_(F.cross_entropy(lg, target))
        if dlogits is not None:
            p = torch.softmax(lg, 1)
            p[torch.arange(len(target)), target] -= 1
            _put(dlogits, p * (grad_scale / len(target)))
        if correct is not None:
            correct.copy_((lg.argmax(1) == target).sum().to(correct.dtype))

    def linear_bwd(self, x, w, dlogits, dw, db, dx):
        self._log("linear_bwd")
        xf = x.float().reshape(x.shape[0], -1)
        _put(dw, dlogits.t() @ xf)                                     # the C entry point clears, then accumulates
        _put(db, dlogits.sum(0))
        _put(dx, dlogits @ w)

    def head_fwd_bwd(self, x, w, b, target, logits, loss, correct, dlogits, dx, dw, db, grad_scale=1.0):
        self._log("head_fwd_bwd")
        xf = x.float().reshape(x.shape[0], -1)
        lg = xf @ w.t() + b
        _put(logits, lg)
        loss.copy_(F.cross_entropy(lg, target))
        correct.copy_((lg.argmax(1) == target).sum().to(correct.dtype))
        p = torch.softmax(lg, 1)
        p[torch.arange(len(target)), target] -= 1
        dl = p * (grad_scale / len(target))
        if dlogits is not None:
            _put(dlogits, dl)
        _put(dx, dl @ w)
        dw += (dl.t() @ xf).reshape(dw.shape)                          # accumulated onto, as documented
        db += dl.sum(0)

    # ---- weight gradients ---------------------------------------------------------------------------
    def conv_wgrad(self, dy, x, dw, taps=9, cin_valid=None, split_k=0, accumulate=False, block_n=0):
        self._log("conv_wgrad")
        B, H, W, Cout = dy.shape
        Cin = x.shape[3]
        cv = Cin if cin_valid is None else cin_valid
        g = dy.float().reshape(-1, Cout)
        if taps == 1:
            full = g.t() @ x.float().reshape(-1, Cin)                    # [Cout][Cin]
            dw += full[:, :cv].reshape(dw.shape)
            return dw
        xp = F.pad(x.float(), (0, 0, 1, 1, 1, 1))
        parts = []
        for r in range(3):
            for s in range(3):
                parts.append(g.t() @ xp[:, r:r + H, s:s + W, :].reshape(-1, Cin))
        full = torch.stack(parts, 1)                                     # [Cout][9][Cin]
        dw += full[:, :, :cv].reshape(dw.shape)
        return dw

    def conv_wgrad_push(self, dy, x, dw, *, counters, world, rank, peer_stage, flat_off, own_start, own_len, stage_base,
                        split_k=0, block_n=0):
        self._log("conv_wgrad_push")
        self.calls.pop()
        out = self.conv_wgrad(dy, x, dw)
        self.calls[-1] = "conv_wgrad_push"
        return out

    def conv_dgrad_bn_reduce(self, dy, wd, dout_prev, out_prev, gamma, beta, sums, block_n=0, split_k=1, workspace=None,
                             cta_pair=False):
        self.conv_igemm(dy, wd, None, dout_prev)
        self.calls[-1] = "conv_dgrad_bn_reduce"
        C = dout_prev.shape[-1]
        g = dout_prev.float().reshape(-1, C)
        a = out_prev.float().reshape(-1, C)
        d = g * (a > 0)
        ig = torch.where(gamma.abs() > 1e-20, 1.0 / gamma, torch.zeros_like(gamma))
        sums[:C] += d.sum(0)
        sums[C:2 * C] += (d * (a - beta) * ig).sum(0)
        return dout_prev

    # ---- optimizer and operand copies ---------------------------------------------------------------
    def sgd_momentum(self, p, g, buf, lr, momentum, wd, grad_scale=1.0):
        self._log("sgd_momentum")
        d = g * grad_scale + wd * p
        buf.mul_(momentum).add_(d)
        p.sub_(lr * buf)

    def BatchedWeightPrep(self, layers):
        layers = list(layers)

        def run():
            self._log("prep_conv_weights_batched")
            for master, wf, wd, cout, cin, taps, kpad in layers:
                w = master.float().reshape(cout, -1)                     # KRSC rows: [Cout][taps*Cin]
                full = torch.zeros(cout, kpad)
                full[:, : w.shape[1]] = w
                _put(wf, full)
                if wd is not None:
                    k = w.view(cout, 3, 3, cin)
                    _put(wd, k.flip(1, 2).permute(3, 1, 2, 0))           # [Cin][r'][s'][Cout]
        return run
