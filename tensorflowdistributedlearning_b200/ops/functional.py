"""Autograd functions that bind the native sm_100a kernels, with the torch oracle as CPU path.

Dispatch rule (one rule, no backend matrix): a CUDA bf16 NHWC tensor whose channel counts meet
the kernels' 16-byte alignment goes to the native library; everything else (CPU towers used by
the plumbing tests, fp32 debugging, odd channel counts such as the 3-channel stem) runs the
plain-torch restatement in ``ops.reference``.  On a GPU the native library is mandatory --
``_native.lib()`` raises if it cannot be loaded, there is no silent fallback.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.nn.functional as F

from . import reference as R

_FORCE_REFERENCE = False  # flipped by tests / the proxy baseline to route everything through torch ops
_LAUNCHES = 0             # native kernel launches issued through this module (bench "gpu_launches")


def force_reference(flag: bool) -> None:
    global _FORCE_REFERENCE
    _FORCE_REFERENCE = bool(flag)


def launches() -> int:
    return _LAUNCHES


def _count(n: int = 1) -> None:
    global _LAUNCHES
    _LAUNCHES += n


def _native_ok(x: torch.Tensor) -> bool:
    return (not _FORCE_REFERENCE) and x.is_cuda and x.dtype == torch.bfloat16


# ---------------------------------------------------------------------------------------------
# convolution
# ---------------------------------------------------------------------------------------------
def conv_pads(h: int, w: int, kh: int, kw: int, stride: int, dil: int, padding) -> Tuple[int, int, int, int, int, int]:
    """-> (P, Q, pad_top, pad_bottom, pad_left, pad_right)"""
    if padding == "SAME":
        P, pt, pb = R.same_pad(h, kh, stride, dil)
        Q, pl, pr = R.same_pad(w, kw, stride, dil)
    elif padding == "VALID":
        pt = pb = pl = pr = 0
        P = (h - ((kh - 1) * dil + 1)) // stride + 1
        Q = (w - ((kw - 1) * dil + 1)) // stride + 1
    else:
        pt, pb, pl, pr = padding
        P = (h + pt + pb - ((kh - 1) * dil + 1)) // stride + 1
        Q = (w + pl + pr - ((kw - 1) * dil + 1)) // stride + 1
    return P, Q, pt, pb, pl, pr


class _ConvNative(torch.autograd.Function):
    """tcgen05 implicit-GEMM forward; dgrad = same kernel on flipped/transposed weights; wgrad = MN-major kernel."""

    @staticmethod
    def forward(ctx, x, weight, bias, stride, dil, pads, out_hw, want_stats=False):
        from . import native_ops as O

        K, Rk, Sk, C = weight.shape
        taps = Rk * Sk
        wf = torch.empty((K, taps, C), dtype=torch.bfloat16, device=x.device)
        need_dgrad = ctx.needs_input_grad[0]
        wd = torch.empty((C, taps, K), dtype=torch.bfloat16, device=x.device) if (need_dgrad and stride == 1) else None
        O.conv2d_weight_prep(weight.detach(), wf, wd, K, taps, C)
        pt, pb, pl, pr = pads
        partials = None
        if want_stats:
            y, partials, nparts = O.conv2d_fprop_stats(x, wf, R=Rk, S=Sk, stride=stride, dil=dil, pad_h=pt, pad_w=pl,
                                                       out_hw=out_hw, bias=None if bias is None else bias.detach())
            partials = partials[:nparts]
        else:
            y = O.conv2d_fprop(x, wf, R=Rk, S=Sk, stride=stride, dil=dil, pad_h=pt, pad_w=pl, out_hw=out_hw,
                               bias=None if bias is None else bias.detach())
        _count(2)
        ctx.save_for_backward(x, wd if wd is not None else wf)
        ctx.cfg = (K, Rk, Sk, C, stride, dil, pads, tuple(x.shape), bias is not None, wd is not None)
        if want_stats:
            ctx.mark_non_differentiable(partials)
            return y, partials
        return y

    @staticmethod
    def backward(ctx, dy, *_unused):
        from . import native_ops as O

        x, wsaved = ctx.saved_tensors
        K, Rk, Sk, C, stride, dil, pads, xshape, has_bias, has_wd = ctx.cfg
        pt, pb, pl, pr = pads
        n, h, w, _ = xshape
        dy = dy.contiguous()
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            if has_wd:
                dx = O.conv2d_fprop(dy, wsaved, R=Rk, S=Sk, stride=1, dil=dil, pad_h=(Rk - 1) * dil - pt,
                                    pad_w=(Sk - 1) * dil - pl, out_hw=(h, w))
                _count(1)
            else:
                # strided data gradient: one tcgen05 launch per input-parity class, strided TMA stores
                dx = O.conv2d_dgrad_strided(dy, wsaved, (n, h, w, C), R=Rk, S=Sk, stride=stride, pad_h=pt, pad_w=pl)
                _count(stride * stride)
        if ctx.needs_input_grad[1]:
            dw = torch.zeros((K, Rk * Sk, C), dtype=torch.float32, device=x.device)
            O.conv2d_wgrad(x, dy, dw, R=Rk, S=Sk, stride=stride, dil=dil, pad_h=pt, pad_w=pl)
            _count(1)
            dw = dw.view(K, Rk, Sk, C)
        if has_bias and ctx.needs_input_grad[2]:
            db = dy.float().sum((0, 1, 2))
        return dx, dw, db, None, None, None, None, None


def conv2d(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None, stride: int = 1,
           dil: int = 1, padding="SAME", want_stats: bool = False):
    """NHWC convolution, weight [K, R, S, C] (fp32 master).  TF 'SAME' geometry by default.

    ``want_stats=True`` returns ``(y, stats)`` where ``stats`` is either None (torch path) or the per-CTA partial
    BatchNorm statistics the tcgen05 kernel accumulated in its epilogue; pass it to :func:`batch_norm`."""
    K, Rk, Sk, C = weight.shape
    n, h, w, c = x.shape
    assert c == C, f"channel mismatch: input {c} vs weight {C}"
    P, Q, pt, pb, pl, pr = conv_pads(h, w, Rk, Sk, stride, dil, padding)
    if _native_ok(x) and C % 8 == 0 and K % 8 == 0 and stride <= 2:
        if want_stats:
            return _ConvNative.apply(x.contiguous(), weight, bias, stride, dil, (pt, pb, pl, pr), (P, Q), True)
        return _ConvNative.apply(x.contiguous(), weight, bias, stride, dil, (pt, pb, pl, pr), (P, Q))
    # torch path (cuDNN on GPU through a channels-last view; plain CPU conv on CPU towers)
    wq = weight.to(x.dtype)
    bq = None if bias is None else bias.to(x.dtype)
    xin = x.permute(0, 3, 1, 2)
    if pt == pb and pl == pr:
        y = F.conv2d(xin, wq.permute(0, 3, 1, 2), bq, stride=stride, padding=(pt, pl), dilation=dil)
    else:
        y = F.conv2d(F.pad(xin, (pl, pr, pt, pb)), wq.permute(0, 3, 1, 2), bq, stride=stride, dilation=dil)
    y = y.permute(0, 2, 3, 1)
    return (y, None) if want_stats else y


# ---------------------------------------------------------------------------------------------
# batch norm (+ residual) (+ ReLU)
# ---------------------------------------------------------------------------------------------
class _BNNative(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, running_mean, running_var, residual, decay, eps, relu, training, partials=None):
        from . import native_ops as O

        shape = x.shape
        C = shape[-1]
        x2 = x.reshape(-1, C)
        r2 = None if residual is None else residual.reshape(-1, C)
        if training and partials is not None:
            mean, invstd, scale, shift = O.bn_finalize_partials(partials, partials.shape[0], x2.shape[0], gamma, beta,
                                                                running_mean, running_var, decay, eps)
            _count(1)
        elif training:
            mean, invstd, scale, shift = O.bn_fwd_stats(x2, gamma, beta, running_mean, running_var, decay, eps)
            _count(2)
        else:
            scale, shift = O.bn_infer_scale_shift(gamma, beta, running_mean, running_var, eps)
            mean = invstd = None
            _count(1)
        z = O.bn_apply(x2, scale, shift, r2, relu)
        _count(1)
        if training:
            need_z = relu and residual is not None
            ctx.save_for_backward(x2, z if need_z else None, scale, shift, mean, invstd)
        ctx.cfg = (relu, residual is not None, training, shape)
        return z.view(shape)

    @staticmethod
    def backward(ctx, dz):
        from . import native_ops as O

        relu, has_res, training, shape = ctx.cfg
        if not training:
            raise RuntimeError("backward through inference-mode BatchNorm is not supported")
        x2, z2, scale, shift, mean, invstd = ctx.saved_tensors
        C = shape[-1]
        dz2 = dz.contiguous().view(-1, C)
        dx, dres, dgamma, dbeta = O.bn_bwd(dz2, x2, z2, scale, shift, mean, invstd, relu, has_res)
        _count(4)
        return (dx.view(shape), dgamma if ctx.needs_input_grad[1] else None,
                dbeta if ctx.needs_input_grad[2] else None, None, None,
                dres.view(shape) if has_res else None, None, None, None, None, None)


def batch_norm(x: torch.Tensor, gamma, beta, running_mean, running_var, *, decay: float, eps: float,
               training: bool, relu: bool = False, residual: Optional[torch.Tensor] = None,
               stats: Optional[torch.Tensor] = None) -> torch.Tensor:
    """y = relu?(BN(x) + residual?) over the last (channel) axis with TF fused-BN semantics.
    ``stats``: partial statistics of ``x`` from ``conv2d(..., want_stats=True)`` (skips the statistics pass)."""
    C = x.shape[-1]
    if _native_ok(x) and C % 8 == 0 and gamma is not None and beta is not None:
        res = None if residual is None else residual.contiguous()
        return _BNNative.apply(x.contiguous(), gamma, beta, running_mean, running_var, res, decay, eps, relu,
                               training, stats if training else None)
    if training:
        y = R.batch_norm_train(x, gamma, beta, running_mean, running_var, decay, eps)
    else:
        y = R.batch_norm_infer(x, gamma, beta, running_mean, running_var, eps)
    if residual is not None:
        y = y + residual
    return F.relu(y) if relu else y


# ---------------------------------------------------------------------------------------------
# losses (native on CUDA, oracle elsewhere)
# ---------------------------------------------------------------------------------------------
class _LovaszNative(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits2d, labels2d):
        from . import native_ops as O

        B = logits2d.shape[0]
        loss, dl = O.lovasz_hinge(logits2d, labels2d, True, 1.0 / B)
        _count(1)
        ctx.save_for_backward(dl)
        return loss.mean()

    @staticmethod
    def backward(ctx, g):
        (dl,) = ctx.saved_tensors
        return dl * g, None


def lovasz_loss(y_true: torch.Tensor, y_pred: torch.Tensor) -> torch.Tensor:
    """Mean over images of the per-image Lovasz hinge of logits ``y_pred`` (reference: core/losses.py:83-92)."""
    B = y_pred.shape[0]
    lg = y_pred.reshape(B, -1).float()
    lb = y_true.reshape(B, -1).float()
    if lg.is_cuda and not _FORCE_REFERENCE and lg.shape[1] <= 16384:
        return _LovaszNative.apply(lg, lb)
    return R.lovasz_hinge_per_image(lg, lb).mean()


class _SoftmaxCENative(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels):
        from . import native_ops as O

        B = logits.shape[0]
        loss, dl, t1, t5 = O.softmax_ce(logits, labels, True, 1.0 / B, 5)
        _count(1)
        ctx.save_for_backward(dl)
        ctx.in_dtype = logits.dtype
        ctx.mark_non_differentiable(t1, t5)
        return loss.mean(), t1, t5

    @staticmethod
    def backward(ctx, g, _g1, _g5):
        (dl,) = ctx.saved_tensors
        return (dl * g).to(ctx.in_dtype), None


def softmax_cross_entropy(logits: torch.Tensor, labels: torch.Tensor):
    """-> (mean loss, top1[B], top5[B]).  Fused log-softmax + NLL + gradient + in-top-k on the GPU."""
    if logits.is_cuda and not _FORCE_REFERENCE:
        return _SoftmaxCENative.apply(logits, labels)
    lf = logits.float()
    loss = F.cross_entropy(lf, labels)
    return loss, R.topk_correct(lf.detach(), labels, 1), R.topk_correct(lf.detach(), labels, min(5, lf.shape[1]))


def seg_metrics(logits: torch.Tensor, labels: torch.Tensor, threshold: float = 0.5, reference_compat: bool = True):
    """Per-image (iou_metric, pixel_accuracy)."""
    import math

    B = logits.shape[0]
    lg = logits.detach().reshape(B, -1)
    lb = labels.detach().reshape(B, -1)
    if lg.is_cuda and not _FORCE_REFERENCE:
        from . import native_ops as O

        thr = math.log(threshold / (1.0 - threshold)) if 0.0 < threshold < 1.0 else 0.0
        _count(1)
        return O.seg_metrics(lg, lb, thr, reference_compat)
    return R.seg_metrics(lg, lb, threshold, reference_compat)
