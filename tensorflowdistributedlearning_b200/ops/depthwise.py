"""Depthwise convolution (NHWC): native CUDA-core kernels on the GPU, torch oracle elsewhere.

reference call sites: core/layers.py:34-42 (ASPP, dilated, bias+ReLU) and
core/xception.py:102-110 (Xception separable convs, stride 1/2).
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn.functional as F

from . import functional as TF
from . import reference as R


class _DepthwiseNative(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, stride, dil, pads, out_hw, relu):
        from . import native_ops as O

        y = O.depthwise_fwd(x, weight.detach(), None if bias is None else bias.detach(), stride, dil, pads[0], pads[2],
                            out_hw, relu)
        TF._count(1)
        ctx.save_for_backward(x, weight, y if relu else None)
        ctx.cfg = (stride, dil, pads, relu, bias is not None)
        return y

    @staticmethod
    def backward(ctx, dy):
        from . import native_ops as O

        x, weight, y = ctx.saved_tensors
        stride, dil, pads, relu, has_bias = ctx.cfg
        dy = dy.contiguous()
        dx, dw, db = O.depthwise_bwd(x, weight.detach(), dy, y, stride, dil, pads[0], pads[2], relu)
        TF._count(2)
        return dx, dw, (db if has_bias else None), None, None, None, None, None


def depthwise_conv2d(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None, stride: int = 1,
                     dil: int = 1, padding="SAME", relu: bool = False) -> torch.Tensor:
    """x [N,H,W,C], weight [kh,kw,C] -> [N,P,Q,C]; optional fused bias + ReLU."""
    kh, kw, C = weight.shape
    n, h, w, c = x.shape
    P, Q, pt, pb, pl, pr = TF.conv_pads(h, w, kh, kw, stride, dil, padding)
    if TF._native_ok(x) and C % 8 == 0:
        return _DepthwiseNative.apply(x.contiguous(), weight, bias, stride, dil, (pt, pb, pl, pr), (P, Q), relu)
    y = R.depthwise_conv2d_nhwc(x, weight.to(x.dtype), stride, dil, (pt, pb, pl, pr), bias)
    return F.relu(y) if relu else y
