"""Pooling ops (NHWC): 3x3/s2 SAME max-pool and global average pool.

reference call sites: core/resnet.py:241 (max_pool2d SAME), :246-248 and :464-465 (reduce_mean).
"""
from __future__ import annotations

import torch

from . import functional as TF
from . import reference as R


class _MaxPoolNative(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, k, stride, pads, out_hw, want_stats=False):
        from . import native_ops as O

        ctx.set_materialize_grads(False)
        ctx.cfg = (k, stride, pads, tuple(x.shape))
        TF._count(1)
        if want_stats:  # + the batch statistics of the pooled tensor for the BatchNorm that follows
            y, idx, partials = O.maxpool_fwd_stats(x, pads[0], pads[2], out_hw)
            ctx.save_for_backward(idx)
            ctx.mark_non_differentiable(partials)
            return y, partials
        y, idx = O.maxpool_fwd(x, k, stride, pads[0], pads[2], out_hw)
        ctx.save_for_backward(idx)
        return y

    @staticmethod
    def backward(ctx, dy, *_unused):
        from . import native_ops as O

        (idx,) = ctx.saved_tensors
        k, stride, pads, in_shape = ctx.cfg
        dx = O.maxpool_bwd(dy.contiguous(), idx, in_shape, k, stride, pads[0], pads[2])
        TF._count(1)
        return dx, None, None, None, None, None


def max_pool_same(x: torch.Tensor, k: int = 3, stride: int = 2, want_stats: bool = False):
    """TF 'SAME' max-pool.  ``want_stats=True`` -> (y, stats): ``stats`` is None (torch path / unsupported shape) or the
    partial BatchNorm statistics of ``y`` accumulated by the pooling kernel itself; pass it to the BatchNorm that
    consumes ``y`` (the reference's pool1 -> postnorm, core/resnet.py:241-242)."""
    n, h, w, c = x.shape
    P, Q, pt, pb, pl, pr = TF.conv_pads(h, w, k, k, stride, 1, "SAME")
    if TF._native_ok(x) and c % 8 == 0:
        fused = (want_stats and k == 3 and stride == 2 and 256 % (c // 8) == 0 and n <= 65535
                 and P * Q * c < (1 << 31))
        if fused:
            return _MaxPoolNative.apply(x.contiguous(), k, stride, (pt, pb, pl, pr), (P, Q), True)
        y = _MaxPoolNative.apply(x.contiguous(), k, stride, (pt, pb, pl, pr), (P, Q))
        return (y, None) if want_stats else y
    y = R.max_pool_same_nhwc(x, k, stride)
    return (y, None) if want_stats else y


class _GapNative(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        from . import native_ops as O

        ctx.shape = x.shape
        TF._count(1)
        return O.gap_fwd(x)

    @staticmethod
    def backward(ctx, dy):
        from . import native_ops as O

        TF._count(1)
        return O.gap_bwd(dy.contiguous(), ctx.shape)


class _GapConv1x1(torch.autograd.Function):
    """Global average pool folded into the 1x1 convolution that follows it (core/resnet.py:464-466)."""

    @staticmethod
    def forward(ctx, x, weight):
        from . import native_ops as O

        K = weight.shape[0]
        w2 = weight.detach().view(K, -1)
        y, gap = O.gap_conv1x1_fwd(x, w2)
        TF._count(1)
        ctx.save_for_backward(gap)
        ctx.weight_param, ctx.in_shape = weight, tuple(x.shape)
        return y

    @staticmethod
    def backward(ctx, dy):
        from . import native_ops as O

        (gap,) = ctx.saved_tensors
        wp = ctx.weight_param
        K = wp.shape[0]
        dy2 = dy.contiguous().view(-1, K)
        direct = TF._direct_ok(wp)
        dw = None
        if ctx.needs_input_grad[1]:
            dw = wp.grad.view(K, -1) if direct else torch.zeros((K, gap.shape[1]), dtype=torch.float32, device=dy.device)
        dx = O.gap_conv1x1_bwd(dy2, wp.detach().view(K, -1), gap, ctx.in_shape, ctx.needs_input_grad[0], dw)
        TF._count(2)
        if direct and dw is not None:
            TF._notify(wp)
            dw = None
        return dx, (None if dw is None else dw.view(wp.shape))


def gap_conv1x1(x: torch.Tensor, weight: torch.Tensor) -> torch.Tensor:
    """[N,H,W,C] -> [N,1,1,K]: conv1x1(global_avg_pool(x)) with ``weight`` [K,1,1,C] (fp32 master, no bias), one kernel
    on the GPU when the shape suits the fused kernel (few images x few output channels: the seg-net's pooling branch)."""
    n, h, w, c = x.shape
    K = weight.shape[0]
    if (TF._native_ok(x) and c % 8 == 0 and c <= 12288 and tuple(weight.shape[1:3]) == (1, 1) and weight.shape[3] == c
            and weight.is_contiguous() and n * ((K + 63) // 64) <= 4096 and n <= 256):
        return _GapConv1x1.apply(x.contiguous(), weight)
    return TF.conv2d(global_avg_pool(x, keepdims=True), weight)


def global_avg_pool(x: torch.Tensor, keepdims: bool = True) -> torch.Tensor:
    """mean over H, W of NHWC -> [N,1,1,C] (keepdims) or [N,C]."""
    n, h, w, c = x.shape
    if TF._native_ok(x) and c % 8 == 0:
        y = _GapNative.apply(x.contiguous())
    else:
        y = x.float().mean((1, 2)).to(x.dtype)
    return y.view(n, 1, 1, c) if keepdims else y


def max_pool(x: torch.Tensor, k: int, stride: int, pads) -> torch.Tensor:
    """Max-pool with explicit (top, bottom, left, right) padding (padded cells never win)."""
    n, h, w, c = x.shape
    pt, pb, pl, pr = pads
    P = (h + pt + pb - k) // stride + 1
    Q = (w + pl + pr - k) // stride + 1
    if TF._native_ok(x) and c % 8 == 0:
        return _MaxPoolNative.apply(x.contiguous(), k, stride, (pt, pb, pl, pr), (P, Q))
    import torch.nn.functional as F

    xin = F.pad(x.permute(0, 3, 1, 2), (pl, pr, pt, pb), value=float("-inf"))
    return F.max_pool2d(xin, k, stride).permute(0, 2, 3, 1)
