"""Plain-PyTorch oracles (CPU or GPU, any float dtype) for every op the native library implements.

These restate the semantics of the TensorFlow 1.x ops the reference calls;
the reference itself cannot run on this stack, so these *are* the numerical spec.  They double as
the bring-up / CPU-tower code path of ``ops.functional``.
"""
from __future__ import annotations

import math
from typing import Optional, Sequence, Tuple

import torch
import torch.nn.functional as F


# ---------------------------------------------------------------------------------------------
# geometry
# ---------------------------------------------------------------------------------------------
def same_pad(in_size: int, k: int, stride: int, dil: int = 1) -> Tuple[int, int, int]:
    """TF padding='SAME' -> (out, pad_before, pad_after); the odd pixel goes after (A1)."""
    out = -(-in_size // stride)
    k_eff = (k - 1) * dil + 1
    total = max((out - 1) * stride + k_eff - in_size, 0)
    return out, total // 2, total - total // 2


def fixed_padding(x: torch.Tensor, kernel_size: int, rate: int = 1, mode: str = "constant") -> torch.Tensor:
    """Explicit (k_eff-1) padding split beg=(k_eff-1)//2, end=rest on H and W of an NHWC tensor.

    reference: core/layers.py:52-79 and core/xception.py:18-35.
    """
    k_eff = kernel_size + (kernel_size - 1) * (rate - 1)
    total = k_eff - 1
    beg = total // 2
    end = total - beg
    return F.pad(x.permute(0, 3, 1, 2), (beg, end, beg, end), mode=mode).permute(0, 2, 3, 1)


# ---------------------------------------------------------------------------------------------
# conv / pool (NHWC, TF semantics)
# ---------------------------------------------------------------------------------------------
def conv2d_nhwc(x: torch.Tensor, w_ohwi: torch.Tensor, stride: int = 1, dil: int = 1, padding="SAME",
                bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """x [N,H,W,C]; w [K,R,S,C].  padding: 'SAME' | 'VALID' | (top, bottom, left, right)."""
    K, R, S, C = w_ohwi.shape
    n, h, wd, c = x.shape
    if padding == "SAME":
        _, pt, pb = same_pad(h, R, stride, dil)
        _, pl, pr = same_pad(wd, S, stride, dil)
    elif padding == "VALID":
        pt = pb = pl = pr = 0
    else:
        pt, pb, pl, pr = padding
    xin = x.permute(0, 3, 1, 2)
    if pt or pb or pl or pr:
        xin = F.pad(xin, (pl, pr, pt, pb))
    y = F.conv2d(xin, w_ohwi.permute(0, 3, 1, 2).to(x.dtype), None if bias is None else bias.to(x.dtype),
                 stride=stride, dilation=dil)
    return y.permute(0, 2, 3, 1)


def depthwise_conv2d_nhwc(x: torch.Tensor, w_rsc: torch.Tensor, stride: int = 1, dil: int = 1, padding="SAME",
                          bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """x [N,H,W,C]; w [R,S,C] (depth multiplier 1)."""
    R, S, C = w_rsc.shape
    n, h, wd, c = x.shape
    if padding == "SAME":
        _, pt, pb = same_pad(h, R, stride, dil)
        _, pl, pr = same_pad(wd, S, stride, dil)
    elif padding == "VALID":
        pt = pb = pl = pr = 0
    else:
        pt, pb, pl, pr = padding
    xin = F.pad(x.permute(0, 3, 1, 2), (pl, pr, pt, pb))
    w = w_rsc.permute(2, 0, 1).unsqueeze(1).to(x.dtype)
    y = F.conv2d(xin, w, None if bias is None else bias.to(x.dtype), stride=stride, dilation=dil, groups=C)
    return y.permute(0, 2, 3, 1)


def max_pool_same_nhwc(x: torch.Tensor, k: int = 3, stride: int = 2) -> torch.Tensor:
    """TF max_pool2d padding='SAME': padded cells never win (-inf)."""
    n, h, w, c = x.shape
    _, pt, pb = same_pad(h, k, stride)
    _, pl, pr = same_pad(w, k, stride)
    xin = F.pad(x.permute(0, 3, 1, 2), (pl, pr, pt, pb), value=float("-inf"))
    return F.max_pool2d(xin, k, stride).permute(0, 2, 3, 1)


def subsample(x: torch.Tensor, factor: int) -> torch.Tensor:
    """resnet_utils.subsample == 1x1 max-pool with stride == strided slice from index 0 (A6)."""
    return x if factor == 1 else x[:, ::factor, ::factor, :]


# ---------------------------------------------------------------------------------------------
# batch norm (TF fused semantics, A5)
# ---------------------------------------------------------------------------------------------
def batch_norm_train(x: torch.Tensor, gamma, beta, running_mean, running_var, decay: float, eps: float,
                     update: bool = True):
    """x [..., C].  Returns y; updates running stats in place (biased var to normalise, unbiased to track)."""
    xf = x.float()
    dims = tuple(range(x.dim() - 1))
    n = xf.numel() // xf.shape[-1]
    mean = xf.mean(dims)
    var = xf.var(dims, unbiased=False)
    y = (xf - mean) * torch.rsqrt(var + eps)
    if gamma is not None:
        y = y * gamma
    if beta is not None:
        y = y + beta
    if update and running_mean is not None:
        with torch.no_grad():
            running_mean.mul_(decay).add_((1 - decay) * mean.detach())
            unb = var.detach() * (n / max(n - 1, 1))
            running_var.mul_(decay).add_((1 - decay) * unb)
    return y.to(x.dtype)


def batch_norm_infer(x: torch.Tensor, gamma, beta, running_mean, running_var, eps: float):
    scale = torch.rsqrt(running_var + eps)
    if gamma is not None:
        scale = scale * gamma
    shift = -running_mean * scale
    if beta is not None:
        shift = shift + beta
    return (x.float() * scale + shift).to(x.dtype)


# ---------------------------------------------------------------------------------------------
# TF1 legacy bilinear resize + the reference's `_upsample` (A9, A10; core/layers.py:82-109)
# ---------------------------------------------------------------------------------------------
def _legacy_axis(in_size: int, out_size: int, device):
    scale = in_size / out_size
    src = torch.arange(out_size, device=device, dtype=torch.float64) * scale
    lo = src.floor().long().clamp_(max=in_size - 1)
    hi = (lo + 1).clamp_(max=in_size - 1)
    w = (src - lo.double())
    return lo, hi, w


def resize_bilinear_legacy(x: torch.Tensor, out_hw: Tuple[int, int]) -> torch.Tensor:
    """tf.image.resize_bilinear(align_corners=False) with TF1 coordinates src = dst * in/out.  x NHWC."""
    n, h, w, c = x.shape
    oh, ow = out_hw
    ylo, yhi, wy = _legacy_axis(h, oh, x.device)
    xlo, xhi, wx = _legacy_axis(w, ow, x.device)
    xf = x.float()
    wy = wy.to(xf.dtype).view(1, oh, 1, 1)
    wx = wx.to(xf.dtype).view(1, 1, ow, 1)
    top = xf[:, ylo][:, :, xlo] * (1 - wx) + xf[:, ylo][:, :, xhi] * wx
    bot = xf[:, yhi][:, :, xlo] * (1 - wx) + xf[:, yhi][:, :, xhi] * wx
    return (top * (1 - wy) + bot * wy).to(x.dtype)


def symmetric_pad1(x: torch.Tensor) -> torch.Tensor:
    """tf.pad(mode='SYMMETRIC') by one pixel on H and W of NHWC (edge pixel repeated)."""
    return F.pad(x.permute(0, 3, 1, 2).float(), (1, 1, 1, 1), mode="replicate").permute(0, 2, 3, 1).to(x.dtype)


def upsample_reference(x: torch.Tensor, out_hw: Tuple[int, int]) -> torch.Tensor:
    """reference `_upsample`: symmetric-pad 1 -> legacy bilinear to (H+4, W+4) -> crop 2 px each side."""
    oh, ow = out_hw
    y = resize_bilinear_legacy(symmetric_pad1(x), (oh + 4, ow + 4))
    return y[:, 2:-2, 2:-2, :]


# ---------------------------------------------------------------------------------------------
# losses
# ---------------------------------------------------------------------------------------------
def lovasz_grad(gt_sorted: torch.Tensor) -> torch.Tensor:
    """reference: core/losses.py:5-15."""
    gts = gt_sorted.sum()
    inter = gts - gt_sorted.cumsum(0)
    union = gts + (1 - gt_sorted).cumsum(0)
    jac = 1.0 - inter / union
    if jac.numel() > 1:
        jac = torch.cat([jac[:1], jac[1:] - jac[:-1]])
    return jac


def lovasz_hinge_flat(logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
    """reference: core/losses.py:40-65 (full descending sort, stable on ties; grad only via relu(errors))."""
    if logits.numel() == 0:
        return logits.sum() * 0.0
    signs = 2.0 * labels - 1.0
    errors = 1.0 - logits * signs
    errors_sorted, perm = torch.sort(errors, descending=True, stable=True)
    gt_sorted = labels[perm]
    grad = lovasz_grad(gt_sorted).detach()
    return torch.dot(F.relu(errors_sorted), grad)


def lovasz_hinge_per_image(logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
    """[B, P] -> [B] per-image losses (mean over B gives reference lovasz_hinge(per_image=True))."""
    return torch.stack([lovasz_hinge_flat(lg, lb.to(lg.dtype)) for lg, lb in zip(logits, labels)])


def lovasz_loss(y_true: torch.Tensor, y_pred: torch.Tensor) -> torch.Tensor:
    """reference: core/losses.py:83-92.  y_* are [B,H,W,1] or [B,H,W]; logits in, scalar out."""
    B = y_pred.shape[0]
    return lovasz_hinge_per_image(y_pred.reshape(B, -1).float(), y_true.reshape(B, -1).float()).mean()


def softmax_cross_entropy(logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
    return F.cross_entropy(logits.float(), labels, reduction="none")


# ---------------------------------------------------------------------------------------------
# metrics
# ---------------------------------------------------------------------------------------------
IOU_THRESHOLDS = [0.5 + 0.05 * k for k in range(10)]  # reference: core/metric.py:3


def seg_metrics(logits: torch.Tensor, labels: torch.Tensor, threshold: float = 0.5, reference_compat: bool = True):
    """Per-image (iou_metric, pixel_accuracy) from logits.  reference: core/metric.py:14-37,60.

    reference_compat=True reproduces ``mean_t(score * [score > t])`` (defect 14); False gives the
    Kaggle metric ``mean_t([score > t])``.
    """
    B = logits.shape[0]
    lg = logits.reshape(B, -1).float()
    y = labels.reshape(B, -1).float() > 0.5
    thr = math.log(threshold / (1.0 - threshold)) if 0.0 < threshold < 1.0 else 0.0
    pred = lg > thr
    tp = (pred & y).sum(1).double()
    fp = (pred & ~y).sum(1).double()
    fn = (~pred & y).sum(1).double()
    den = tp + fp + fn
    score = torch.where(den == 0, torch.ones_like(den), tp / den.clamp(min=1))
    t = torch.tensor(IOU_THRESHOLDS, dtype=torch.float64, device=lg.device)
    hit = (score[:, None] > t[None, :]).double()
    if reference_compat:
        hit = hit * score[:, None]
    iou = hit.mean(1).float()
    acc = (pred == y).float().mean(1)
    return iou, acc


def topk_correct(logits: torch.Tensor, labels: torch.Tensor, k: int) -> torch.Tensor:
    """[B] float: 1 where the label is within the k largest logits (ties -> lower index wins, A21)."""
    xl = logits.gather(1, labels[:, None])
    idx = torch.arange(logits.shape[1], device=logits.device)[None, :]
    rank = ((logits > xl) | ((logits == xl) & (idx < labels[:, None]))).sum(1)
    return (rank < k).float()


# ---------------------------------------------------------------------------------------------
# optimizers (A19, A20)
# ---------------------------------------------------------------------------------------------
def exponential_decay(lr0: float, step: int, decay_steps: float, decay_rate: float) -> float:
    return lr0 if decay_steps <= 0 else lr0 * decay_rate ** (step / decay_steps)


def adam_tf_step(p, g, m, v, lr: float, t: int, beta1=0.9, beta2=0.999, eps=1e-8):
    lr_t = lr * math.sqrt(1 - beta2 ** t) / (1 - beta1 ** t)
    m.mul_(beta1).add_((1 - beta1) * g)
    v.mul_(beta2).add_((1 - beta2) * g * g)
    p.sub_(lr_t * m / (v.sqrt() + eps))


def sgd_momentum_step(p, g, mom, lr: float, momentum: float, weight_decay: float = 0.0):
    g = g + weight_decay * p
    mom.mul_(momentum).add_(g)
    p.sub_(lr * mom)


# ---------------------------------------------------------------------------------------------
# initialisers (A24)
# ---------------------------------------------------------------------------------------------
def variance_scaling_(w_ohwi: torch.Tensor, generator: Optional[torch.Generator] = None) -> torch.Tensor:
    """contrib variance_scaling_initializer(): truncated normal, stddev = sqrt(1.3 * 2 / fan_in)."""
    K, R, S, C = w_ohwi.shape
    std = math.sqrt(1.3 * 2.0 / (R * S * C))
    return truncated_normal_(w_ohwi, std, generator)


def truncated_normal_(t: torch.Tensor, std: float, generator: Optional[torch.Generator] = None) -> torch.Tensor:
    with torch.no_grad():
        torch.nn.init.trunc_normal_(t, mean=0.0, std=std, a=-2 * std, b=2 * std, generator=generator)
    return t
