"""Depthwise convolution (NHWC): native CUDA-core kernels on the GPU, torch oracle elsewhere.

reference call sites: core/layers.py:34-42 (ASPP, dilated, bias+ReLU) and
core/xception.py:102-110 (Xception separable convs, stride 1/2).

3x3 filters (every use in the reference) take the strip kernels of ``csrc/depthwise.cu``: the pre-activation ReLU of
the Xception module can be folded into the load (``relu_in``) and the ASPP ReLU into the store (``relu``); the data
gradient of a stride-1 layer is the same kernel on flipped taps.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn.functional as F

from . import functional as TF
from . import reference as R


class _DepthwiseNative(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, stride, dil, pads, out_hw, relu, relu_in, want_stats=False, want_tap=False):
        from . import native_ops as O

        ctx.want_tap = bool(want_tap)

        kh, kw, C = weight.shape
        fast = kh == 3 and kw == 3
        wd = weight.detach().float().contiguous()
        bd = None if bias is None else bias.detach().float().contiguous()
        partials = None
        if fast and want_stats:
            y, partials = O.depthwise3x3_fwd(x, wd.view(9, C), bd, None, stride, dil, pads[0], pads[2], out_hw, relu_in,
                                             relu, want_stats=True)
        elif fast:
            y = O.depthwise3x3_fwd(x, wd.view(9, C), bd, None, stride, dil, pads[0], pads[2], out_hw, relu_in, relu)
        else:
            xin = torch.relu(x) if relu_in else x
            y = O.depthwise_fwd(xin, wd, bd, stride, dil, pads[0], pads[2], out_hw, relu)
        TF._count(1)
        ctx.save_for_backward(x, weight, y if relu else None)
        ctx.weight_param, ctx.bias_param = weight, bias
        ctx.set_materialize_grads(False)  # no zero-fill for the (non-differentiable) statistics output
        ctx.cfg = (stride, dil, pads, relu, relu_in, bias is not None, fast)
        outs = [y]
        if want_stats:
            if partials is None:  # shape not served by the fused kernel: an empty marker the caller maps to "no stats"
                partials = torch.empty(0, device=x.device)
            ctx.mark_non_differentiable(partials)
            outs.append(partials)
        if want_tap:
            # an alias of the input for the enclosing module's skip connection: its gradient comes back to THIS node
            # and is added inside the data-gradient kernel (after the input-ReLU mask) instead of by aten::add
            outs.append(x.view_as(x))
        return outs[0] if len(outs) == 1 else tuple(outs)

    @staticmethod
    def backward(ctx, dy, *rest):
        from . import native_ops as O

        x, weight, y = ctx.saved_tensors
        d_tap = rest[-1] if (ctx.want_tap and rest and rest[-1] is not None) else None
        if dy is None:
            return (d_tap,) + (None,) * 10
        stride, dil, pads, relu, relu_in, has_bias, fast = ctx.cfg
        dy = dy.contiguous()
        kh, kw, C = weight.shape
        wd = weight.detach().float().contiguous()
        if fast:
            wp, bp = ctx.weight_param, ctx.bias_param
            ymask = y if relu else None
            if TF._direct_ok(wp) and (not has_bias or TF._direct_ok(bp)):
                # straight into the flat gradient bucket, and -- nothing downstream in backward reads it -- on the
                # weight-gradient side stream (see _ConvNative.backward)
                def run():
                    O.depthwise3x3_wgrad(x, dy, ymask, stride, dil, pads[0], pads[2], relu_in, has_bias,
                                         dw_out=wp.grad, db_out=bp.grad if has_bias else None)
                if TF._ASYNC_WGRAD:
                    side = TF._side_stream()
                    side.wait_stream(torch.cuda.current_stream())
                    with torch.cuda.stream(side):
                        run()
                    for t in (x, dy, ymask):
                        if t is not None:
                            t.record_stream(side)
                    TF._SIDE_DIRTY = True
                else:
                    run()
                TF._notify(wp)
                if has_bias:
                    TF._notify(bp)
                dw = db = None
            else:
                dw, db = O.depthwise3x3_wgrad(x, dy, ymask, stride, dil, pads[0], pads[2], relu_in, has_bias)
                dw = dw.view(3, 3, C)
            TF._count(1)
            dx = None
            if ctx.needs_input_grad[0]:
                if relu and (stride == 1 or (stride == 2 and dil == 1)):
                    # backward of the fused output ReLU (reference split_separable_conv2d: depthwise + bias + ReLU,
                    # core/layers.py:6-49), then the same fast data-gradient kernels as the no-activation case
                    dy = torch.ops.aten.threshold_backward(dy, y, 0)
                    relu = False
                    TF._count(1)
                if stride == 1 and not relu:
                    n, h, w, _ = x.shape
                    dx = O.depthwise3x3_dgrad_s1(dy, wd.contiguous().view(9, C), x if relu_in else None, dil,
                                                 2 * dil - pads[0], 2 * dil - pads[2],
                                                 addend=None if d_tap is None else d_tap.contiguous())
                    d_tap = None
                elif stride == 2 and dil == 1 and not relu:
                    dx = O.depthwise3x3_dgrad_s2(dy, wd.view(9, C), x if relu_in else None, tuple(x.shape), pads[0],
                                                 pads[2])
                else:
                    dx, _, _ = O.depthwise_bwd(x, wd, dy, y, stride, dil, pads[0], pads[2], relu)
                    if relu_in:
                        dx = dx * (x > 0)
                TF._count(1)
        else:
            xin = torch.relu(x) if relu_in else x
            dx, dw, db = O.depthwise_bwd(xin, wd, dy, y, stride, dil, pads[0], pads[2], relu)
            if relu_in:
                dx = dx * (x > 0)
            TF._count(2)
        if d_tap is not None:
            dx = d_tap if dx is None else dx + d_tap
        return dx, dw, (db if has_bias else None), None, None, None, None, None, None, None, None


def depthwise_conv2d(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None, stride: int = 1,
                     dil: int = 1, padding="SAME", relu: bool = False, relu_in: bool = False, want_stats: bool = False,
                     want_tap: bool = False):
    """x [N,H,W,C], weight [kh,kw,C] -> [N,P,Q,C]; optional ReLU on the input, fused bias + ReLU on the output.

    ``want_stats=True`` returns ``(y, stats)``: ``stats`` are the partial BatchNorm statistics of ``y`` accumulated by the
    depthwise kernel itself (pass them to :func:`ops.functional.batch_norm`), or None when they were not produced."""
    kh, kw, C = weight.shape
    n, h, w, c = x.shape
    P, Q, pt, pb, pl, pr = TF.conv_pads(h, w, kh, kw, stride, dil, padding)
    if want_tap:
        # (y, stats_or_None, x_alias): feed x_alias to whatever else reads x (a skip connection) and that gradient is
        # added inside this layer's data-gradient kernel
        if TF._native_ok(x) and C % 8 == 0 and torch.is_grad_enabled():
            y, st, tap = _DepthwiseNative.apply(x.contiguous(), weight, bias, stride, dil, (pt, pb, pl, pr), (P, Q),
                                                relu, relu_in, True, True)
            return y, (st if st.numel() else None), tap
        out = depthwise_conv2d(x, weight, bias, stride, dil, padding, relu, relu_in, True)
        return out[0], out[1], x
    if TF._native_ok(x) and C % 8 == 0:
        if want_stats:
            y, st = _DepthwiseNative.apply(x.contiguous(), weight, bias, stride, dil, (pt, pb, pl, pr), (P, Q), relu,
                                           relu_in, True)
            return y, (st if st.numel() else None)
        return _DepthwiseNative.apply(x.contiguous(), weight, bias, stride, dil, (pt, pb, pl, pr), (P, Q), relu, relu_in)
    xin = F.relu(x) if relu_in else x
    y = R.depthwise_conv2d_nhwc(xin, weight.to(x.dtype), stride, dil, (pt, pb, pl, pr), bias)
    y = F.relu(y) if relu else y
    return (y, None) if want_stats else y
