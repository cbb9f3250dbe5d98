"""The reference's `_upsample` (symmetric pad 1 -> TF1 legacy bilinear -> crop 2): one fused gather kernel.

reference: core/layers.py:82-109.  TF1's resize_bilinear(align_corners=False) uses src = dst * in/out
(no half-pixel centres), which differs from torch.nn.functional.interpolate -- hence its own kernel/oracle.
"""
from __future__ import annotations

from typing import Tuple

import torch

from . import functional as TF
from . import reference as R


class _UpsampleNative(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, out_hw):
        from . import native_ops as O

        ctx.in_shape = x.shape
        ctx.out_hw = out_hw
        TF._count(1)
        return O.upsample_fwd(x, out_hw)

    @staticmethod
    def backward(ctx, dy):
        from . import native_ops as O

        TF._count(1)
        return O.upsample_bwd(dy.contiguous(), ctx.in_shape, ctx.out_hw), None


def upsample(x: torch.Tensor, out_hw: Tuple[int, int]) -> torch.Tensor:
    """x NHWC -> [N, out_h, out_w, C] with the reference's exact interpolation weights."""
    if x.is_cuda and not TF._FORCE_REFERENCE and x.dtype in (torch.bfloat16, torch.float32):
        return _UpsampleNative.apply(x.contiguous(), tuple(out_hw))
    return R.upsample_reference(x, tuple(out_hw))
