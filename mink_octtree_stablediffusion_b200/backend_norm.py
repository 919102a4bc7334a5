"""Fused normalisation on the engine's two passes (csrc/norm.cu): one statistics
read (`mk_seg_sums`) and one normalise + affine read/write (`mk_seg_affine`).

`segment_norm` is the one autograd Function behind MinkowskiBatchNorm /
MinkowskiSyncBatchNorm and the instance-norm family (reference:
MinkowskiNormalization.py:55-195, :198-622).  The layer-specific part — how the
per-(segment, channel) mean and variance turn into the affine coefficients
y = A x + C (batch-norm weight/bias, the fork's group folding, AdaIN scale/shift)
— is a small torch function on [n_seg, C] tensors and is differentiated by
autograd itself; only the two passes over [N, C] are kernels:

    forward   S = seg_sums(x - K, (x - K)^2)      mean = K + S0/m,  var = S1/m - (S0/m)^2
              A, C = coef(mean, var, *params)     y = A[seg] x + C[seg]
    backward  dA = sum dy x, dC = sum dy          (one seg_sums pass)
              dmean, dvar, dparams = autograd(coef)
              dx = A dy + (2 dvar / m) (x - mean) + dmean / m      (one affine pass)
"""
from __future__ import annotations

import ctypes
from typing import Callable, Optional, Sequence

import torch

from . import _lib
from .backend import _dtype_code, _ptr, _stream


def _acc_dtype(t: torch.Tensor) -> torch.dtype:
    return torch.float64 if t.dtype == torch.float64 else torch.float32


def seg_sums(a: torch.Tensor, b: torch.Tensor, shift_a: Optional[torch.Tensor], shift_b: Optional[torch.Tensor],
             seg: Optional[torch.Tensor], n_seg: int) -> torch.Tensor:
    """[n_seg, 2, C]: sum (a - shift_a), sum (a - shift_a) * (b - shift_b) per segment and channel."""
    lib = _lib.load()
    n, c = a.shape
    out = torch.empty((n_seg, 2, c), dtype=_acc_dtype(a), device=a.device)
    _lib.check(lib.mk_seg_sums(_dtype_code(a), _ptr(a), _ptr(b), _ptr(shift_a), _ptr(shift_b), _ptr(seg), n, c, n_seg,
                               ctypes.c_void_p(out.data_ptr()), _stream()))
    return out


def seg_affine(a: torch.Tensor, b: Optional[torch.Tensor], seg: Optional[torch.Tensor], coef_a: torch.Tensor,
               coef_b: Optional[torch.Tensor], coef_c: torch.Tensor) -> torch.Tensor:
    """out[r] = coef_a[seg[r]] * a[r] + coef_b[seg[r]] * b[r] + coef_c[seg[r]]."""
    lib = _lib.load()
    n, c = a.shape
    out = torch.empty_like(a)
    _lib.check(lib.mk_seg_affine(_dtype_code(a), _ptr(a), _ptr(b), _ptr(seg), n, c, _ptr(coef_a), _ptr(coef_b),
                                 _ptr(coef_c), _ptr(out), _stream()))
    return out


class _SegmentNorm(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, seg, n_seg, counts, shift, coef_fn, reduce_fn, stats_out, *params):
        x = x.contiguous()
        acc = _acc_dtype(x)
        S = seg_sums(x, x, shift, shift, seg, n_seg)
        m = counts
        if reduce_fn is not None:            # SyncBatchNorm: sums and counts over all ranks
            S, m = reduce_fn(S, m)
        m = m.to(acc).clamp(min=1).view(-1, 1)
        d = S[:, 0] / m
        mean = d if shift is None else shift + d
        var = (S[:, 1] / m - d * d).clamp(min=0)
        if stats_out is not None:
            stats_out.append((mean.detach(), var.detach(), m))
        mean_l, var_l = mean.detach().requires_grad_(True), var.detach().requires_grad_(True)
        with torch.enable_grad():
            p_l = [p.detach().requires_grad_(p.requires_grad) if isinstance(p, torch.Tensor) else p for p in params]
            A, C = coef_fn(mean_l, var_l, *p_l)
            A, C = A.to(acc).expand(n_seg, x.shape[1]).contiguous(), C.to(acc).expand(n_seg, x.shape[1]).contiguous()
        y = seg_affine(x, None, seg, A.detach(), None, C.detach())
        ctx.save_for_backward(x, seg, mean.detach(), m)
        ctx.graph = (mean_l, var_l, p_l, A, C)
        ctx.n_seg, ctx.reduce_fn = n_seg, reduce_fn
        return y

    @staticmethod
    def backward(ctx, dy):
        x, seg, mean, m = ctx.saved_tensors
        mean_l, var_l, p_l, A, C = ctx.graph
        dy = dy.contiguous()
        if dy.dtype != x.dtype:
            dy = dy.to(x.dtype)
        # dC = sum dy, dA = sum dy * x  (x taken relative to the mean: no cancellation; dA is corrected below)
        S = seg_sums(dy, x, None, mean, seg, ctx.n_seg)
        dC = S[:, 0]
        dA = S[:, 1] + mean * dC
        wanted = [mean_l, var_l] + [p for p in p_l if isinstance(p, torch.Tensor) and p.requires_grad]
        grads = torch.autograd.grad([A, C], wanted, [dA, dC], allow_unused=True)
        dmean = grads[0] if grads[0] is not None else torch.zeros_like(mean)
        dvar = grads[1] if grads[1] is not None else torch.zeros_like(mean)
        if ctx.reduce_fn is not None:        # the statistics were global: so are their gradients
            both, _ = ctx.reduce_fn(torch.stack([dmean, dvar], dim=1), None)
            dmean, dvar = both[:, 0], both[:, 1]
        cb = 2.0 * dvar / m
        cc = dmean / m - cb * mean
        dx = seg_affine(dy, x, seg, A.detach(), cb.contiguous(), cc.contiguous())
        out, gi = [], 2
        for p in p_l:
            if isinstance(p, torch.Tensor) and p.requires_grad:
                out.append(grads[gi])
                gi += 1
            else:
                out.append(None)
        return (dx, None, None, None, None, None, None, None, *out)


def segment_norm(x: torch.Tensor, seg: Optional[torch.Tensor], n_seg: int, counts: torch.Tensor,
                 coef_fn: Callable, params: Sequence = (), shift: Optional[torch.Tensor] = None, reduce_fn=None,
                 stats_out: Optional[list] = None) -> torch.Tensor:
    """y[r] = A[seg[r]] * x[r] + C[seg[r]] with (A, C) = coef_fn(mean, var, *params), where mean / var
    ([n_seg, C], biased variance) are the statistics of x over each segment's rows.

    counts [n_seg]: rows per segment; shift [n_seg, C] (optional): any value near the segment mean, used
    to centre the one-pass sums; reduce_fn(sums, counts) -> (sums, counts) all-reduces them over ranks;
    stats_out (optional list) receives (mean, var, count) for running-statistics updates."""
    return _SegmentNorm.apply(x, seg, n_seg, counts, shift, coef_fn, reduce_fn, stats_out, *params)


# ---------------------------------------------------------------------------
# batch norm: the same two passes with the [C]-sized algebra in one launch each
# ---------------------------------------------------------------------------
def _seg_sums_into(out, a, b, shift_a, shift_b):
    lib = _lib.load()
    n, c = a.shape
    _lib.check(lib.mk_seg_sums(_dtype_code(a), _ptr(a), _ptr(b), _ptr(shift_a), _ptr(shift_b), None, n, c, 1,
                               ctypes.c_void_p(out.data_ptr()), _stream()))


class _BatchNormTrain(torch.autograd.Function):
    """y = batch_norm(x) with batch statistics (torch.nn.BatchNorm1d in training mode,
    MinkowskiNormalization.py:55-103); `reduce_fn` makes the statistics global (SyncBatchNorm)."""

    @staticmethod
    def forward(ctx, x, weight, bias, running_mean, running_var, momentum, eps, reduce_fn):
        lib = _lib.load()
        x = x.contiguous()
        n, c = x.shape
        acc = _acc_dtype(x)
        f64 = int(acc == torch.float64)
        dev = x.device
        buf = torch.empty((7, c), dtype=acc, device=dev)      # sums[2], A, C, mean, invstd, shift
        sums, coef_a, coef_c, mean, invstd, shift = buf[0:2], buf[2], buf[3], buf[4], buf[5], buf[6]
        # parameters / running statistics of a 16-bit module (net.to(bfloat16)): the algebra runs on
        # accumulator-dtype copies, the running statistics are written back in their own dtype
        w_in, rm_in, rv_in = weight, running_mean, running_var
        if weight is not None and weight.dtype != acc:
            weight = weight.detach().to(acc)
        if bias is not None and bias.dtype != acc:
            bias = bias.detach().to(acc)
        if running_mean is not None and running_mean.dtype != acc:
            running_mean, running_var = running_mean.to(acc), running_var.to(acc)
        count_dev = None
        if reduce_fn is None:
            shift.copy_(x[0])                                 # any value near the mean centres the one-pass sums
        elif running_mean is not None:
            shift.copy_(running_mean)                         # ... the same one on every rank
        else:
            shift.zero_()
        _seg_sums_into(sums, x, x, shift, shift)
        if reduce_fn is not None:
            flat = torch.cat([sums.reshape(-1), torch.full((1,), float(n), dtype=acc, device=dev)])
            reduce_fn(flat)
            sums = flat[:2 * c].view(2, c)
            count_dev = flat[2 * c:]
        upd = running_mean is not None and momentum is not None
        _lib.check(lib.mk_bn_coef_fwd(f64, _ptr(sums), _ptr(shift), _ptr(count_dev), float(n), _ptr(weight), _ptr(bias),
                                      float(eps), c, _ptr(coef_a), _ptr(coef_c), _ptr(mean), _ptr(invstd),
                                      _ptr(running_mean) if upd else None, _ptr(running_var) if upd else None,
                                      float(momentum or 0.0), _stream()))
        if upd and rm_in is not running_mean:
            rm_in.copy_(running_mean)
            rv_in.copy_(running_var)
        y = seg_affine(x, None, None, coef_a, None, coef_c)
        ctx.save_for_backward(x, weight, mean, invstd, count_dev if count_dev is not None else torch.empty(0))
        ctx.reduce_fn, ctx.n = reduce_fn, n
        ctx.has_bias = bias is not None
        ctx.param_dtype = w_in.dtype if w_in is not None else (bias.dtype if bias is not None else acc)
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = _lib.load()
        x, weight, mean, invstd, count_dev = ctx.saved_tensors
        dy = dy.contiguous()
        if dy.dtype != x.dtype:
            dy = dy.to(x.dtype)
        n, c = x.shape
        acc = mean.dtype
        buf = torch.empty((7, c), dtype=acc, device=x.device)  # sums[2], cA, cB, cC, dweight, dbias
        sums = buf[0:2]
        _seg_sums_into(sums, dy, x, None, mean)
        if ctx.reduce_fn is not None:
            ctx.reduce_fn(sums)
        _lib.check(lib.mk_bn_coef_bwd(int(acc == torch.float64), _ptr(sums), _ptr(mean), _ptr(invstd),
                                      _ptr(count_dev) if count_dev.numel() else None, float(ctx.n), _ptr(weight), c,
                                      _ptr(buf[2]), _ptr(buf[3]), _ptr(buf[4]), _ptr(buf[5]), _ptr(buf[6]), _stream()))
        dx = seg_affine(dy, x, None, buf[2], buf[3], buf[4])
        # under SyncBatchNorm sums are global; the parameter gradients a rank reports are its LOCAL share in
        # torch's implementation, and DDP sums them - the global value divided by the world size keeps that sum
        dw, db = buf[5], buf[6]
        if ctx.reduce_fn is not None:
            import torch.distributed as dist
            w = dist.get_world_size()
            dw, db = dw / w, db / w
        if ctx.param_dtype != acc:
            dw, db = dw.to(ctx.param_dtype), db.to(ctx.param_dtype)
        return (dx, dw if weight is not None else None, db if ctx.has_bias else None, None, None, None, None, None)


def batch_norm_train(x, weight, bias, running_mean, running_var, momentum, eps, reduce_fn=None):
    """weight / bias / running statistics: any floating dtype (the algebra runs in float32, or float64 for
    float64 features); reduce_fn(flat_tensor) all-reduces in place."""
    return _BatchNormTrain.apply(x, weight, bias, running_mean, running_var, momentum, eps, reduce_fn)


# ---------------------------------------------------------------------------
# instance / group norm: the two passes with the [n_seg, C] algebra in one launch each way
# ---------------------------------------------------------------------------
class _GroupInstanceNorm(torch.autograd.Function):
    """Per batch item, statistics over the item's rows and `gs` adjacent channels; weight / bias
    shared by `gw` adjacent channels (MinkowskiInstanceNorm gs = gw = 1, MinkowskiStableInstanceNorm
    gs = gw = group, MinkowskiStableGroupNorm gs = C, gw = 1; reference MinkowskiNormalization.py:198-478).
    The same result as `segment_norm` with the layers' coefficient functions (tested against it), with
    six launches per norm and pass instead of ~100."""

    @staticmethod
    def forward(ctx, x, seg, n_seg, counts, shift, weight, bias, eps, gs, gw):
        lib = _lib.load()
        x = x.contiguous()
        n, c = x.shape
        acc = _acc_dtype(x)
        f64 = int(acc == torch.float64)
        S = seg_sums(x, x, shift, shift, seg, n_seg)
        w = weight.detach().reshape(-1).to(acc).contiguous() if weight is not None else None
        b = bias.detach().reshape(-1).to(acc).contiguous() if bias is not None else None
        buf = torch.empty((3, n_seg, c), dtype=acc, device=x.device)          # A, C, mean
        stat = torch.empty((2, n_seg, c // gs), dtype=acc, device=x.device)   # mu, inv
        _lib.check(lib.mk_gn_coef_fwd(f64, _ptr(S), _ptr(shift), _ptr(counts), _ptr(w), _ptr(b), float(eps), c, gs, gw,
                                      n_seg, _ptr(buf[0]), _ptr(buf[1]), _ptr(buf[2]), _ptr(stat[0]), _ptr(stat[1]),
                                      _stream()))
        y = seg_affine(x, None, seg, buf[0], None, buf[1])
        ctx.save_for_backward(x, seg, buf, stat, counts, w if w is not None else torch.empty(0))
        ctx.meta = (n_seg, gs, gw, weight.shape if weight is not None else None, weight.dtype if weight is not None else None,
                    bias.shape if bias is not None else None, bias.dtype if bias is not None else None)
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = _lib.load()
        x, seg, buf, stat, counts, w = ctx.saved_tensors
        n_seg, gs, gw, w_shape, w_dtype, b_shape, b_dtype = ctx.meta
        dy = dy.contiguous()
        if dy.dtype != x.dtype:
            dy = dy.to(x.dtype)
        n, c = x.shape
        acc = buf.dtype
        mean = buf[2]
        S = seg_sums(dy, x, None, mean, seg, n_seg)
        co = torch.empty((3, n_seg, c), dtype=acc, device=x.device)
        need_w, need_b = w_shape is not None and ctx.needs_input_grad[5], b_shape is not None and ctx.needs_input_grad[6]
        dwb = torch.empty((2, c // gw), dtype=acc, device=x.device) if (need_w or need_b) else None
        _lib.check(lib.mk_gn_coef_bwd(int(acc == torch.float64), _ptr(S), _ptr(mean), _ptr(stat[0]), _ptr(stat[1]),
                                      _ptr(counts), _ptr(w) if w.numel() else None, c, gs, gw, n_seg, _ptr(co[0]),
                                      _ptr(co[1]), _ptr(co[2]), _ptr(dwb[0]) if need_w else None,
                                      _ptr(dwb[1]) if need_b else None, _stream()))
        dx = seg_affine(dy, x, seg, co[0], co[1], co[2]) if ctx.needs_input_grad[0] else None
        dw = dwb[0].to(w_dtype).view(w_shape) if need_w else None
        db = dwb[1].to(b_dtype).view(b_shape) if need_b else None
        return (dx, None, None, None, None, dw, db, None, None, None)


def group_instance_norm(x, seg, n_seg, counts, shift, weight, bias, eps, gs, gw):
    """counts / shift: accumulator dtype (float32, or float64 for float64 features)."""
    return _GroupInstanceNorm.apply(x, seg, n_seg, counts, shift, weight, bias, eps, gs, gw)
