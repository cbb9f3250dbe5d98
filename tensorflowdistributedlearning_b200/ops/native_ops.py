"""Thin tensor-level wrappers over the C ABI of ``libtdl_b200.so``.

Each function validates shapes/dtypes, allocates outputs with torch (so CUDA-graph capture and
the caching allocator keep working) and launches on the *current* torch stream.  Nothing here
falls back to a torch implementation: if the library or the GPU is missing the call raises.
The autograd glue lives in ``ops/functional.py``; the plain-torch oracles in ``ops/reference.py``.
"""
from __future__ import annotations

import ctypes
from typing import Optional, Tuple

import torch

from .. import _native as N

_i = ctypes.c_int
_i64 = ctypes.c_int64
_f = ctypes.c_float


def _chk_bf16_nhwc(t: torch.Tensor, name: str):
    if t.dtype != torch.bfloat16 or not t.is_cuda or not t.is_contiguous():
        raise ValueError(f"{name} must be a contiguous CUDA bf16 tensor, got {t.dtype} {t.device} contiguous={t.is_contiguous()}")


def same_pad(in_size: int, k: int, stride: int, dil: int = 1) -> Tuple[int, int, int]:
    """TF 'SAME' geometry: (out, pad_before, pad_after).  """
    out = -(-in_size // stride)
    k_eff = (k - 1) * dil + 1
    total = max((out - 1) * stride + k_eff - in_size, 0)
    return out, total // 2, total - total // 2


# --------------------------------------------------------------------------------------------
# convolution (tcgen05)
# --------------------------------------------------------------------------------------------
def conv2d_fprop(x: torch.Tensor, w: torch.Tensor, *, R: int, S: int, stride: int = 1, dil: int = 1,
                 pad_h: int = 0, pad_w: int = 0, out_hw: Optional[Tuple[int, int]] = None,
                 bias: Optional[torch.Tensor] = None, residual: Optional[torch.Tensor] = None,
                 relu: bool = False, out_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
    """y[N,P,Q,K] = conv(x[N,H,W,C], w[K, R*S, C]) with fused bias / residual / ReLU epilogue.

    ``out_mask`` ([N*P*Q, K/8] uint8, the 1-bit ReLU mask written by ``bn_apply``; needs ``residual``, no bias/ReLU):
    outputs whose bit is clear are zeroed -- the ReLU backward of the tensor this data gradient belongs to."""
    _chk_bf16_nhwc(x, "x")
    _chk_bf16_nhwc(w, "w")
    n, h, wd, c = x.shape
    k = w.shape[0]
    assert w.numel() == k * R * S * c, (w.shape, k, R, S, c)
    if out_hw is None:
        p = (h + 2 * pad_h - ((R - 1) * dil + 1)) // stride + 1
        q = (wd + 2 * pad_w - ((S - 1) * dil + 1)) // stride + 1
    else:
        p, q = out_hw
    y = torch.empty((n, p, q, k), dtype=torch.bfloat16, device=x.device)
    if residual is not None:
        _chk_bf16_nhwc(residual, "residual")
        assert residual.shape == y.shape
    if bias is not None:
        assert bias.dtype == torch.float32 and bias.numel() == k and bias.is_contiguous()
    if out_mask is not None:
        assert residual is not None and bias is None and not relu and out_mask.dtype == torch.uint8
        assert out_mask.numel() == n * p * q * (k // 8) and k % 32 == 0
        rc = N.lib().tdl_conv2d_fprop_res_mask_bf16(N.ptr(x), N.ptr(w), N.ptr(y), N.ptr(residual), N.ptr(out_mask),
                                                    _i(n), _i(h), _i(wd), _i(c), _i(k), _i(R), _i(S), _i(stride),
                                                    _i(dil), _i(pad_h), _i(pad_w), _i(p), _i(q), N.stream_ptr())
        N.check(rc, "tdl_conv2d_fprop_res_mask_bf16")
        return y
    rc = N.lib().tdl_conv2d_fprop_bf16(N.ptr(x), N.ptr(w), N.ptr(y), N.ptr(bias), N.ptr(residual),
                                       _i(n), _i(h), _i(wd), _i(c), _i(k), _i(R), _i(S), _i(stride), _i(dil),
                                       _i(pad_h), _i(pad_w), _i(p), _i(q), _i(int(relu)), N.stream_ptr())
    N.check(rc, "tdl_conv2d_fprop_bf16")
    return y


def conv2d_dgrad_bnstats(dy: torch.Tensor, w: torch.Tensor, *, R: int, S: int, dil: int = 1, pad_h: int = 0,
                         pad_w: int = 0, out_hw, bn_x: torch.Tensor, bn_mean: torch.Tensor, bn_invstd: torch.Tensor,
                         residual: Optional[torch.Tensor] = None, out_mask: Optional[torch.Tensor] = None,
                         bn_scale: Optional[torch.Tensor] = None, bn_shift: Optional[torch.Tensor] = None):
    """Stride-1 data gradient dx = conv(dy, w) (+ residual) (* mask) that also reduces, in its epilogue, the sums
    BatchNorm backward needs for the BatchNorm that produced this tensor (``bn_x`` = that BatchNorm's input).
    Returns (dx, partials [nparts, 2, K]) with partials = per-CTA (sum dz, sum dz * xhat).  With ``bn_scale`` /
    ``bn_shift`` the sums see the recomputed ReLU mask ``x * scale + shift > 0`` while dx is stored unmasked."""
    _chk_bf16_nhwc(dy, "dy")
    _chk_bf16_nhwc(w, "w")
    n, h, wd, c = dy.shape
    k = w.shape[0]
    p, q = out_hw
    assert w.numel() == k * R * S * c and bn_x.is_contiguous() and bn_x.numel() == n * p * q * k
    assert bn_x.dtype == torch.bfloat16 and bn_mean.dtype == torch.float32 and bn_invstd.dtype == torch.float32
    dx = torch.empty((n, p, q, k), dtype=torch.bfloat16, device=dy.device)
    if residual is not None:
        _chk_bf16_nhwc(residual, "residual")
        assert residual.shape == dx.shape
    if out_mask is not None:
        assert out_mask.dtype == torch.uint8 and out_mask.numel() == n * p * q * (k // 8) and k % 32 == 0
    partials = torch.empty((int(N.lib().tdl_num_sms()), 2, k), dtype=torch.float32, device=dy.device)
    nparts = ctypes.c_int(0)
    rc = N.lib().tdl_conv2d_dgrad_bnstats_bf16(N.ptr(dy), N.ptr(w), N.ptr(dx), N.ptr(residual), N.ptr(out_mask),
                                               N.ptr(bn_x), N.ptr(bn_mean), N.ptr(bn_invstd), N.ptr(bn_scale),
                                               N.ptr(bn_shift), _i(n), _i(h), _i(wd),
                                               _i(c), _i(k), _i(R), _i(S), _i(dil), _i(pad_h), _i(pad_w), _i(p), _i(q),
                                               N.ptr(partials), ctypes.byref(nparts), N.stream_ptr())
    N.check(rc, "tdl_conv2d_dgrad_bnstats_bf16")
    return dx, partials[:int(nparts.value)]


def conv2d_fprop_stats(x: torch.Tensor, w: torch.Tensor, *, R: int, S: int, stride: int = 1, dil: int = 1,
                       pad_h: int = 0, pad_w: int = 0, out_hw=None, bias: Optional[torch.Tensor] = None,
                       residual: Optional[torch.Tensor] = None, relu: bool = False):
    """Convolution + per-CTA partial BatchNorm statistics of its (bf16) output in the same kernel.
    With ``residual`` (and optionally ``relu``): statistics of ``relu?(conv + bias + residual)`` as stored.
    Returns (y, partials [sms, 2, K] fp32, nparts)."""
    _chk_bf16_nhwc(x, "x")
    _chk_bf16_nhwc(w, "w")
    n, h, wd, c = x.shape
    k = w.shape[0]
    p, q = out_hw
    y = torch.empty((n, p, q, k), dtype=torch.bfloat16, device=x.device)
    partials = torch.empty((int(N.lib().tdl_num_sms()), 2, k), dtype=torch.float32, device=x.device)
    nparts = ctypes.c_int(0)
    if residual is not None:
        _chk_bf16_nhwc(residual, "residual")
        assert residual.shape == y.shape
        rc = N.lib().tdl_conv2d_fprop_res_stats_bf16(N.ptr(x), N.ptr(w), N.ptr(y), N.ptr(bias), N.ptr(residual), _i(n),
                                                     _i(h), _i(wd), _i(c), _i(k), _i(R), _i(S), _i(stride), _i(dil),
                                                     _i(pad_h), _i(pad_w), _i(p), _i(q), _i(int(relu)),
                                                     N.ptr(partials), ctypes.byref(nparts), N.stream_ptr())
        N.check(rc, "tdl_conv2d_fprop_res_stats_bf16")
        return y, partials, int(nparts.value)
    assert not relu, "ReLU with statistics needs the residual form"
    rc = N.lib().tdl_conv2d_fprop_stats_bf16(N.ptr(x), N.ptr(w), N.ptr(y), N.ptr(bias), _i(n), _i(h), _i(wd), _i(c),
                                             _i(k), _i(R), _i(S), _i(stride), _i(dil), _i(pad_h), _i(pad_w), _i(p),
                                             _i(q), N.ptr(partials), ctypes.byref(nparts), N.stream_ptr())
    N.check(rc, "tdl_conv2d_fprop_stats_bf16")
    return y, partials, int(nparts.value)


def bn_finalize_partials(partials: torch.Tensor, nparts: int, rows: int, gamma, beta, running_mean, running_var,
                         decay: float, eps: float, mean_offset: Optional[torch.Tensor] = None):
    """``mean_offset`` [C]: a constant per channel that is missing from the tensor the partial sums describe (a folded
    convolution bias); it is added to the batch mean for the moving-mean update only."""
    C = partials.shape[-1]
    out = torch.empty((4, C), dtype=torch.float32, device=partials.device)
    rc = N.lib().tdl_bn_finalize_partials_off(N.ptr(partials), _i(nparts), _i64(rows), _i(C), N.ptr(gamma),
                                              N.ptr(beta), N.ptr(running_mean), N.ptr(running_var), _f(decay), _f(eps),
                                              N.ptr(out[0]), N.ptr(out[1]), N.ptr(out[2]), N.ptr(out[3]),
                                              N.ptr(mean_offset), N.stream_ptr())
    N.check(rc, "tdl_bn_finalize_partials")
    return out[0], out[1], out[2], out[3]


def conv2d_dgrad_strided(dy: torch.Tensor, wd: torch.Tensor, in_shape, *, R: int, S: int, stride: int, pad_h: int,
                         pad_w: int, zero_buffers: Optional[dict] = None, zero_key=None) -> torch.Tensor:
    """Data gradient of a stride-`stride` convolution (dilation 1) on the tcgen05 kernel.

    dx is split into stride*stride input-parity classes; class (a, b) is a stride-1 convolution of dy with the taps
    r = r0 + stride*i, s = s0 + stride*j, written through a strided TMA store.  ``wd``: the flipped / transposed
    data-gradient copy [C, R*S (flipped), K] of the weights (``conv2d_weight_prep``); every class reads its taps out of
    it through a (base, row stride, column stride) walk, so no per-class weight copy is made.

    Classes without a tap (1x1 stride 2: three of four) must read as zero.  ``zero_buffers`` (a dict owned by the
    caller, e.g. the trainer's weight-prep registry) keeps one pre-zeroed dx per (``zero_key`` = the layer, shape): the
    kernels only ever write the non-empty classes, so the zeros survive from step to step and the per-step fill of the
    whole tensor goes away.  The returned tensor is then SHARED between steps: consume it before the layer's next
    backward.  Without a dict / key the tensor is zero-filled on every call."""
    _chk_bf16_nhwc(dy, "dy")
    n, h, w, c = in_shape
    _, p, q, k = dy.shape
    assert wd.dtype == torch.bfloat16 and wd.is_contiguous() and wd.numel() == c * R * S * k
    empty_class = False
    classes = []
    for a in range(stride):
        r0 = (a + pad_h) % stride
        nr = len(range(r0, R, stride))
        for b in range(stride):
            s0 = (b + pad_w) % stride
            ns = len(range(s0, S, stride))
            ha, wb = -(-(h - a) // stride), -(-(w - b) // stride)
            if ha <= 0 or wb <= 0:
                continue
            if nr == 0 or ns == 0:
                empty_class = True
                continue
            classes.append((a, b, r0, s0, nr, ns, ha, wb))
    if not empty_class:
        dx = torch.empty(in_shape, dtype=torch.bfloat16, device=dy.device)
    else:
        key = (zero_key, tuple(in_shape))
        dx = zero_buffers.get(key) if (zero_buffers is not None and zero_key is not None) else None
        if dx is None:
            dx = torch.zeros(in_shape, dtype=torch.bfloat16, device=dy.device)
            if zero_buffers is not None and zero_key is not None and not torch.cuda.is_current_stream_capturing():
                zero_buffers[key] = dx
    for (a, b, r0, s0, nr, ns, ha, wb) in classes:
        # class taps, flipped: position (r', s') reads W[k, r0 + stride*(nr-1-r'), s0 + stride*(ns-1-s'), c], which in
        # the fully flipped copy sits at tap ((R-1-r0-stride*(nr-1)) + stride*r') * S + (S-1-s0-stride*(ns-1)) + stride*s'
        base = (R - 1 - r0 - stride * (nr - 1)) * S + (S - 1 - s0 - stride * (ns - 1))
        oh = (a + pad_h - r0) // stride
        ow = (b + pad_w - s0) // stride
        rc = N.lib().tdl_conv2d_fprop_strided_out_taps_bf16(
            N.ptr(dy), N.ptr(wd), N.ptr(dx), _i(n), _i(p), _i(q), _i(k), _i(c), _i(nr), _i(ns), _i(R * S), _i(base),
            _i(stride * S), _i(stride), _i(1), _i(nr - 1 - oh), _i(ns - 1 - ow), _i(ha), _i(wb), _i(h), _i(w),
            _i(stride), _i(a), _i(b), N.stream_ptr())
        N.check(rc, "tdl_conv2d_fprop_strided_out_taps_bf16")
    return dx


def conv2d_wgrad(x: torch.Tensor, dy: torch.Tensor, dw: torch.Tensor, *, R: int, S: int, stride: int = 1,
                 dil: int = 1, pad_h: int = 0, pad_w: int = 0) -> torch.Tensor:
    """dw[K, R*S, C] (fp32) += dy^T (*) x.  ``dw`` is accumulated into (zero it for a fresh gradient)."""
    _chk_bf16_nhwc(x, "x")
    _chk_bf16_nhwc(dy, "dy")
    n, h, wd, c = x.shape
    _, p, q, k = dy.shape
    assert dw.dtype == torch.float32 and dw.is_contiguous() and dw.numel() == k * R * S * c
    rc = N.lib().tdl_conv2d_wgrad_bf16(N.ptr(x), N.ptr(dy), N.ptr(dw), _i(n), _i(h), _i(wd), _i(c), _i(k), _i(R),
                                       _i(S), _i(stride), _i(dil), _i(pad_h), _i(pad_w), _i(p), _i(q), N.stream_ptr())
    N.check(rc, "tdl_conv2d_wgrad_bf16")
    return dw


def conv2d_weight_prep(w_master: torch.Tensor, w_fprop: Optional[torch.Tensor], w_dgrad: Optional[torch.Tensor],
                       K: int, taps: int, C: int) -> None:
    assert w_master.dtype == torch.float32 and w_master.is_contiguous()
    rc = N.lib().tdl_conv2d_weight_prep(N.ptr(w_master), N.ptr(w_fprop), N.ptr(w_dgrad), _i(K), _i(taps), _i(C),
                                        N.stream_ptr())
    N.check(rc, "tdl_conv2d_weight_prep")


# --------------------------------------------------------------------------------------------
# batch norm
# --------------------------------------------------------------------------------------------
_scratch = {}


def _bn_scratch(rows: int, C: int, device) -> torch.Tensor:
    need = int(N.lib().tdl_bn_scratch_floats(_i64(rows), _i(C)))
    if need < 0:
        raise ValueError(f"BN kernels need C % 8 == 0, got {C}")
    key = (device, torch.cuda.current_stream().cuda_stream)
    buf = _scratch.get(key)
    if buf is None or buf.numel() < need:
        buf = torch.empty(max(need, 1 << 20), dtype=torch.float32, device=device)
        _scratch[key] = buf
    return buf


def bn_fwd_stats(x2d: torch.Tensor, gamma, beta, running_mean, running_var, decay: float, eps: float):
    """x2d: [rows, C] bf16.  Returns (mean, invstd, scale, shift) fp32 and updates the moving stats."""
    rows, C = x2d.shape
    dev = x2d.device
    out = torch.empty((4, C), dtype=torch.float32, device=dev)
    scratch = _bn_scratch(rows, C, dev)
    rc = N.lib().tdl_bn_fwd_stats(N.ptr(x2d), _i64(rows), _i(C), N.ptr(scratch), _i64(scratch.numel()), N.ptr(gamma),
                                  N.ptr(beta), N.ptr(running_mean), N.ptr(running_var), _f(decay), _f(eps),
                                  N.ptr(out[0]), N.ptr(out[1]), N.ptr(out[2]), N.ptr(out[3]), N.stream_ptr())
    N.check(rc, "tdl_bn_fwd_stats")
    return out[0], out[1], out[2], out[3]


def bn_infer_scale_shift(gamma, beta, running_mean, running_var, eps: float):
    C = running_mean.numel()
    out = torch.empty((2, C), dtype=torch.float32, device=running_mean.device)
    rc = N.lib().tdl_bn_infer_scale_shift(N.ptr(gamma), N.ptr(beta), N.ptr(running_mean), N.ptr(running_var), _f(eps),
                                          N.ptr(out[0]), N.ptr(out[1]), _i(C), N.stream_ptr())
    N.check(rc, "tdl_bn_infer_scale_shift")
    return out[0], out[1]


def _row_pitch(t2d: torch.Tensor, C: int) -> int:
    """Elements between consecutive rows of a [rows, C] tensor that may be a channel slice of a wider buffer."""
    assert t2d.dim() == 2 and t2d.shape[1] == C and t2d.stride(1) == 1, (t2d.shape, t2d.stride())
    ld = t2d.stride(0) if t2d.shape[0] > 1 else max(t2d.stride(0), C)
    assert ld >= C and ld % 8 == 0 and t2d.data_ptr() % 16 == 0, (ld, C)
    return int(ld)


def bn_apply(x2d, scale, shift, residual2d, relu: bool, want_mask: bool = False, out: Optional[torch.Tensor] = None):
    """y = relu(x*scale + shift (+ residual)).  ``want_mask`` additionally returns the 1-bit-per-element ReLU mask
    ([rows, C/8] uint8) that the backward kernels accept in place of the activation.  ``out``: write into this
    [rows, C] tensor instead of a new one -- it may be a channel slice of a wider buffer (row pitch > C), which is how a
    concatenation of normalised branches costs no copy."""
    rows, C = x2d.shape
    y = torch.empty_like(x2d) if out is None else out
    mask = torch.empty((rows, C // 8), dtype=torch.uint8, device=x2d.device) if (want_mask and relu) else None
    rc = N.lib().tdl_bn_apply_ld(N.ptr(x2d), N.ptr(scale), N.ptr(shift), N.ptr(residual2d), N.ptr(y), N.ptr(mask),
                                 _i64(rows), _i(C), _i(int(relu)), _i64(_row_pitch(y, C)), N.stream_ptr())
    N.check(rc, "tdl_bn_apply_ld")
    return (y, mask) if want_mask else y


def bn_partial_stats(x2d: torch.Tensor):
    """Partial (sum, sum of squares) rows of x2d [rows, C] -> (scratch [nparts, 2, C] view, nparts): the statistics
    pass alone, for ``bn_apply_fused`` when no producing kernel left partials."""
    rows, C = x2d.shape
    scratch = _bn_scratch(rows, C, x2d.device)
    nparts = ctypes.c_int(0)
    rc = N.lib().tdl_bn_partial_stats(N.ptr(x2d), _i64(rows), _i(C), N.ptr(scratch), _i64(scratch.numel()),
                                      ctypes.byref(nparts), N.stream_ptr())
    N.check(rc, "tdl_bn_partial_stats")
    return scratch, int(nparts.value)


def bn_apply_fused(x2d, partials, nparts: int, stat_rows: int, gamma, beta, running_mean, running_var, decay: float,
                   eps: float, residual2d, relu: bool, want_mask: bool = False):
    """Statistics finalize + moving averages + y = relu?(x*scale + shift (+ residual)) in ONE launch (thread-block
    clusters: the first CTA of every cluster folds the partial rows, the others read the constants through DSMEM).
    Returns (y, mask_or_None, mean, invstd, scale, shift)."""
    rows, C = x2d.shape
    y = torch.empty_like(x2d)
    mask = torch.empty((rows, C // 8), dtype=torch.uint8, device=x2d.device) if (want_mask and relu) else None
    out = torch.empty((4, C), dtype=torch.float32, device=x2d.device)
    rc = N.lib().tdl_bn_apply_fused(N.ptr(x2d), N.ptr(partials), _i(nparts), _i64(stat_rows), N.ptr(gamma),
                                    N.ptr(beta), N.ptr(running_mean), N.ptr(running_var), _f(decay), _f(eps),
                                    N.ptr(out[0]), N.ptr(out[1]), N.ptr(out[2]), N.ptr(out[3]), N.ptr(residual2d),
                                    N.ptr(y), N.ptr(mask), _i64(rows), _i(C), _i(int(relu)), N.stream_ptr())
    N.check(rc, "tdl_bn_apply_fused")
    return y, mask, out[0], out[1], out[2], out[3]


def bn_bwd_fused(dz2d, x2d, z2d, scale, shift, mean, invstd, relu: bool, want_dres: bool,
                 acc_dgamma=None, acc_dbeta=None, partials=None):
    """Same contract as ``bn_bwd`` in two launches instead of three: the (mask-specialised) reduction pass, then ONE
    kernel that folds the partial rows into dgamma / dbeta and writes dx (and dres)."""
    rows, C = x2d.shape
    dev = x2d.device
    dgb = torch.empty((2, C), dtype=torch.float32, device=dev)
    zbits = int(z2d is not None and z2d.dtype == torch.uint8)
    if partials is None:
        scratch = _bn_scratch(rows, C, dev)
        nparts = ctypes.c_int(0)
        rc = N.lib().tdl_bn_bwd_partial(N.ptr(dz2d), N.ptr(x2d), N.ptr(z2d), _i(zbits), N.ptr(scale), N.ptr(shift),
                                        N.ptr(mean), N.ptr(invstd), _i64(rows), _i(C), _i(int(relu)), N.ptr(scratch),
                                        _i64(scratch.numel()), ctypes.byref(nparts), N.stream_ptr())
        N.check(rc, "tdl_bn_bwd_partial")
        partials, np_ = scratch, int(nparts.value)
    else:
        assert partials.dtype == torch.float32 and partials.shape[1:] == (2, C) and partials.is_contiguous()
        np_ = partials.shape[0]
    dx = torch.empty_like(x2d)
    dres = torch.empty_like(x2d) if want_dres else None
    rc = N.lib().tdl_bn_bwd_apply_fused(N.ptr(dz2d), N.ptr(x2d), N.ptr(z2d), _i(zbits), N.ptr(scale), N.ptr(shift),
                                        N.ptr(mean), N.ptr(invstd), N.ptr(partials), _i(np_), N.ptr(dgb[0]),
                                        N.ptr(dgb[1]), N.ptr(acc_dgamma), N.ptr(acc_dbeta), N.ptr(dx), N.ptr(dres),
                                        _i64(rows), _i(C), _i(int(relu)), N.stream_ptr())
    N.check(rc, "tdl_bn_bwd_apply_fused")
    return dx, dres, dgb[0], dgb[1]


def bn_bwd(dz2d, x2d, z2d, scale, shift, mean, invstd, relu: bool, want_dres: bool,
           acc_dgamma=None, acc_dbeta=None, partials=None, addend=None, mask_by_x: bool = False):
    """Returns (dx, dres_or_None, dgamma, dbeta).  ``partials`` ([nparts, 2, C], from ``conv2d_dgrad_bnstats``):
    the reduction over (dz, x) has already been done by the kernel that produced dz.  ``addend`` ([rows, C] bf16): a
    second gradient of x, added inside the apply kernel; ``mask_by_x``: x is a ReLU output, dx is zeroed where x == 0."""
    rows, C = x2d.shape
    dev = x2d.device
    dgb = torch.empty((2, C), dtype=torch.float32, device=dev)
    zbits = int(z2d is not None and z2d.dtype == torch.uint8)  # the 1-bit ReLU mask written by bn_apply
    ld = _row_pitch(dz2d, C)  # dz may be a channel slice of a concat buffer's gradient (row pitch > C)
    if partials is not None:
        assert partials.dtype == torch.float32 and partials.shape[1:] == (2, C) and partials.is_contiguous()
        rc = N.lib().tdl_bn_bwd_finalize_partials(N.ptr(partials), _i(partials.shape[0]), _i(C), N.ptr(dgb[0]),
                                                  N.ptr(dgb[1]), N.ptr(acc_dgamma), N.ptr(acc_dbeta), N.stream_ptr())
        N.check(rc, "tdl_bn_bwd_finalize_partials")
    else:
        rc = _bn_bwd_reduce(dz2d, x2d, z2d, zbits, scale, shift, mean, invstd, rows, C, relu, dgb, acc_dgamma,
                            acc_dbeta, ld)
    dx = torch.empty_like(x2d)
    dres = torch.empty_like(x2d) if want_dres else None
    if addend is not None:
        assert addend.dtype == torch.bfloat16 and addend.is_contiguous() and addend.numel() == rows * C
    rc = N.lib().tdl_bn_bwd_apply_ex(N.ptr(dz2d), N.ptr(x2d), N.ptr(z2d), _i(zbits), N.ptr(scale), N.ptr(shift),
                                     N.ptr(mean), N.ptr(invstd), N.ptr(dgb[0]), N.ptr(dgb[1]), N.ptr(dx), N.ptr(dres),
                                     _i64(rows), _i(C), _i(int(relu)), _i64(ld), N.ptr(addend), _i(int(mask_by_x)),
                                     N.stream_ptr())
    N.check(rc, "tdl_bn_bwd_apply_ex")
    return dx, dres, dgb[0], dgb[1]


def _bn_bwd_reduce(dz2d, x2d, z2d, zbits, scale, shift, mean, invstd, rows, C, relu, dgb, acc_dgamma, acc_dbeta,
                   ld=None):
    scratch = _bn_scratch(rows, C, x2d.device)
    ld = C if ld is None else ld
    # the mask-specialised reduction with four rows in flight (128-bit loads, fewer registers than the generic kernel:
    # [802816, 64] 56 -> 48 us, [802816, 256] 185 -> 148 us, profiles/bn_fold_experiment.txt), then the fp64 fold
    nparts = ctypes.c_int(0)
    rc = N.lib().tdl_bn_bwd_partial_ld(N.ptr(dz2d), N.ptr(x2d), N.ptr(z2d), _i(zbits), N.ptr(scale), N.ptr(shift),
                                       N.ptr(mean), N.ptr(invstd), _i64(rows), _i(C), _i(int(relu)), _i64(ld),
                                       N.ptr(scratch), _i64(scratch.numel()), ctypes.byref(nparts), N.stream_ptr())
    N.check(rc, "tdl_bn_bwd_partial_ld")
    rc = N.lib().tdl_bn_bwd_finalize_partials(N.ptr(scratch), _i(int(nparts.value)), _i(C), N.ptr(dgb[0]),
                                              N.ptr(dgb[1]), N.ptr(acc_dgamma), N.ptr(acc_dbeta), N.stream_ptr())
    N.check(rc, "tdl_bn_bwd_finalize_partials")
    return rc


# --------------------------------------------------------------------------------------------
# optimizer
# --------------------------------------------------------------------------------------------
def lr_schedule_step(global_step: torch.Tensor, lr_state: torch.Tensor, base_lr: float, decay_rate: float = 1.0,
                     decay_steps: float = 0.0, beta1: float = 0.9, beta2: float = 0.999) -> None:
    assert global_step.dtype == torch.int64 and lr_state.dtype == torch.float32 and lr_state.numel() >= 2
    rc = N.lib().tdl_lr_schedule_step(N.ptr(global_step), _f(base_lr), _f(decay_rate), _f(decay_steps), _f(beta1),
                                      _f(beta2), N.ptr(lr_state), N.stream_ptr())
    N.check(rc, "tdl_lr_schedule_step")


def sgd_momentum_flat(param, grad, mom, shadow, lr_state, momentum: float, weight_decay: float = 0.0,
                      grad_scale: float = 1.0, zero_grad: bool = True) -> None:
    n = param.numel()
    assert param.dtype == torch.float32 and mom.dtype == torch.float32 and grad.numel() == n
    rc = N.lib().tdl_sgd_momentum_flat(N.ptr(param), N.ptr(grad), N.ptr(mom), N.ptr(shadow), _i64(n), N.ptr(lr_state),
                                       _f(momentum), _f(weight_decay), _f(grad_scale),
                                       _i(int(grad.dtype == torch.bfloat16)), _i(int(zero_grad)), N.stream_ptr())
    N.check(rc, "tdl_sgd_momentum_flat")


def adam_flat(param, grad, m1, m2, shadow, lr_state, beta1: float = 0.9, beta2: float = 0.999, eps: float = 1e-8,
              grad_scale: float = 1.0, zero_grad: bool = True) -> None:
    n = param.numel()
    assert param.dtype == torch.float32 and grad.numel() == n
    rc = N.lib().tdl_adam_flat(N.ptr(param), N.ptr(grad), N.ptr(m1), N.ptr(m2), N.ptr(shadow), _i64(n),
                               N.ptr(lr_state), _f(beta1), _f(beta2), _f(eps), _f(grad_scale),
                               _i(int(grad.dtype == torch.bfloat16)), _i(int(zero_grad)), N.stream_ptr())
    N.check(rc, "tdl_adam_flat")


def cast_f32_to_bf16(src: torch.Tensor, dst: torch.Tensor) -> None:
    rc = N.lib().tdl_cast_f32_to_bf16(N.ptr(src), N.ptr(dst), _i64(src.numel()), N.stream_ptr())
    N.check(rc, "tdl_cast_f32_to_bf16")


# --------------------------------------------------------------------------------------------
# losses / metrics
# --------------------------------------------------------------------------------------------
def lovasz_hinge(logits: torch.Tensor, labels: torch.Tensor, want_grad: bool = True, grad_scale: float = 1.0):
    """logits, labels: [B, P] fp32.  Returns (loss_per_image[B], dlogits[B,P] or None)."""
    B, P = logits.shape
    logits = logits.contiguous().float()
    labels = labels.contiguous().float()
    loss = torch.empty(B, dtype=torch.float32, device=logits.device)
    dl = torch.empty_like(logits) if want_grad else None
    rc = N.lib().tdl_lovasz_hinge(N.ptr(logits), N.ptr(labels), _i(B), _i(P), _f(grad_scale), N.ptr(loss), N.ptr(dl),
                                  N.stream_ptr())
    N.check(rc, "tdl_lovasz_hinge")
    return loss, dl


def seg_metrics(logits: torch.Tensor, labels: torch.Tensor, logit_threshold: float = 0.0,
                reference_compat: bool = True):
    B, P = logits.shape
    logits = logits.contiguous().float()
    labels = labels.contiguous().float()
    out = torch.empty((2, B), dtype=torch.float32, device=logits.device)
    rc = N.lib().tdl_seg_metrics(N.ptr(logits), N.ptr(labels), _i(B), _i(P), _f(logit_threshold),
                                 _i(int(reference_compat)), N.ptr(out[0]), N.ptr(out[1]), N.stream_ptr())
    N.check(rc, "tdl_seg_metrics")
    return out[0], out[1]


def softmax_ce(logits: torch.Tensor, labels: torch.Tensor, want_grad: bool = True, grad_scale: float = 1.0,
               topk: int = 5):
    """logits [B, classes] fp32, labels [B] int64 -> (loss_rows, dlogits|None, top1, topk) all fp32."""
    B, C = logits.shape
    logits = logits.contiguous().float()
    labels = labels.contiguous().to(torch.int64)
    out = torch.empty((3, B), dtype=torch.float32, device=logits.device)
    dl = torch.empty_like(logits) if want_grad else None
    rc = N.lib().tdl_softmax_ce(N.ptr(logits), N.ptr(labels), _i(B), _i(C), _f(grad_scale), _i(topk), N.ptr(out[0]),
                                N.ptr(dl), N.ptr(out[1]), N.ptr(out[2]), N.stream_ptr())
    N.check(rc, "tdl_softmax_ce")
    return out[0], dl, out[1], out[2]


# --------------------------------------------------------------------------------------------
# spatial ops
# --------------------------------------------------------------------------------------------
def maxpool_fwd(x, k, stride, pad_h, pad_w, out_hw):
    _chk_bf16_nhwc(x, "x")
    n, h, w, c = x.shape
    p, q = out_hw
    y = torch.empty((n, p, q, c), dtype=torch.bfloat16, device=x.device)
    idx = torch.empty((n, p, q, c), dtype=torch.uint8, device=x.device)
    rc = N.lib().tdl_maxpool_fwd(N.ptr(x), N.ptr(y), N.ptr(idx), _i(n), _i(h), _i(w), _i(c), _i(p), _i(q), _i(k),
                                 _i(stride), _i(pad_h), _i(pad_w), N.stream_ptr())
    N.check(rc, "tdl_maxpool_fwd")
    return y, idx


def maxpool_fwd_stats(x, pad_h, pad_w, out_hw):
    """3x3 / stride-2 max-pool + per-CTA partial (sum, sum of squares) of the pooled tensor in the same kernel.
    Returns (y, idx, partials [nparts, 2, C])."""
    _chk_bf16_nhwc(x, "x")
    n, h, w, c = x.shape
    p, q = out_hw
    y = torch.empty((n, p, q, c), dtype=torch.bfloat16, device=x.device)
    idx = torch.empty((n, p, q, c), dtype=torch.uint8, device=x.device)
    rows = int(N.lib().tdl_maxpool_stats_rows(_i(n), _i(c), _i(p), _i(q)))
    partials = torch.empty((rows, 2, c), dtype=torch.float32, device=x.device)
    nparts = ctypes.c_int(0)
    rc = N.lib().tdl_maxpool_fwd_stats(N.ptr(x), N.ptr(y), N.ptr(idx), _i(n), _i(h), _i(w), _i(c), _i(p), _i(q),
                                       _i(pad_h), _i(pad_w), N.ptr(partials), ctypes.byref(nparts), N.stream_ptr())
    N.check(rc, "tdl_maxpool_fwd_stats")
    return y, idx, partials[:int(nparts.value)]


def maxpool_bwd(dy, idx, in_shape, k, stride, pad_h, pad_w):
    n, h, w, c = in_shape
    _, p, q, _ = dy.shape
    dx = torch.empty(in_shape, dtype=torch.bfloat16, device=dy.device)
    rc = N.lib().tdl_maxpool_bwd(N.ptr(dy), N.ptr(idx), N.ptr(dx), _i(n), _i(h), _i(w), _i(c), _i(p), _i(q), _i(k),
                                 _i(stride), _i(pad_h), _i(pad_w), N.stream_ptr())
    N.check(rc, "tdl_maxpool_bwd")
    return dx


def gap_conv1x1_fwd(x, w_master):
    """y [N, 1, 1, K] = mean_hw(x [N, H, W, C]) . w [K, C]^T in one kernel; also returns the pooled vectors [N, C] fp32."""
    _chk_bf16_nhwc(x, "x")
    n, h, wd, c = x.shape
    k = w_master.shape[0]
    assert w_master.dtype == torch.float32 and w_master.is_contiguous() and w_master.numel() == k * c
    y = torch.empty((n, 1, 1, k), dtype=torch.bfloat16, device=x.device)
    gap = torch.empty((n, c), dtype=torch.float32, device=x.device)
    rc = N.lib().tdl_gap_conv1x1_fwd(N.ptr(x), N.ptr(w_master), N.ptr(y), N.ptr(gap), _i(n), _i(h * wd), _i(c), _i(k),
                                     N.stream_ptr())
    N.check(rc, "tdl_gap_conv1x1_fwd")
    return y, gap


def gap_conv1x1_bwd(dy, w_master, gap, in_shape, want_dx: bool, dw: Optional[torch.Tensor]):
    """dx [N,H,W,C] bf16 (if ``want_dx``) and ``dw`` [K, C] fp32 += (if given) of the fused pool + 1x1."""
    n, h, wd, c = in_shape
    k = w_master.shape[0]
    dx = torch.empty(in_shape, dtype=torch.bfloat16, device=dy.device) if want_dx else None
    rc = N.lib().tdl_gap_conv1x1_bwd(N.ptr(dy), N.ptr(w_master), N.ptr(gap), N.ptr(dx), N.ptr(dw), _i(n), _i(h * wd),
                                     _i(c), _i(k), N.stream_ptr())
    N.check(rc, "tdl_gap_conv1x1_bwd")
    return dx


def gap_fwd(x):
    _chk_bf16_nhwc(x, "x")
    n, h, w, c = x.shape
    y = torch.empty((n, c), dtype=torch.bfloat16, device=x.device)
    rc = N.lib().tdl_gap_fwd(N.ptr(x), N.ptr(y), _i(n), _i(h * w), _i(c), N.stream_ptr())
    N.check(rc, "tdl_gap_fwd")
    return y


def gap_bwd(dy, in_shape):
    n, h, w, c = in_shape
    dx = torch.empty(in_shape, dtype=torch.bfloat16, device=dy.device)
    rc = N.lib().tdl_gap_bwd(N.ptr(dy), N.ptr(dx), _i(n), _i(h * w), _i(c), N.stream_ptr())
    N.check(rc, "tdl_gap_bwd")
    return dx


def depthwise_fwd(x, w_rsc, bias, stride, dil, pad_h, pad_w, out_hw, relu):
    _chk_bf16_nhwc(x, "x")
    n, h, w, c = x.shape
    r, s, _ = w_rsc.shape
    p, q = out_hw
    wf = w_rsc.float().contiguous()
    bf = None if bias is None else bias.float().contiguous()
    y = torch.empty((n, p, q, c), dtype=torch.bfloat16, device=x.device)
    rc = N.lib().tdl_depthwise_fwd(N.ptr(x), N.ptr(wf), N.ptr(bf), N.ptr(y), _i(n), _i(h), _i(w), _i(c), _i(p), _i(q),
                                   _i(r), _i(s), _i(stride), _i(dil), _i(pad_h), _i(pad_w), _i(int(relu)),
                                   N.stream_ptr())
    N.check(rc, "tdl_depthwise_fwd")
    return y


def depthwise_bwd(x, w_rsc, dy, y, stride, dil, pad_h, pad_w, relu):
    n, h, w, c = x.shape
    r, s, _ = w_rsc.shape
    _, p, q, _ = dy.shape
    wf = w_rsc.float().contiguous()
    dx = torch.empty_like(x)
    dw = torch.zeros((r, s, c), dtype=torch.float32, device=x.device)
    db = torch.zeros((c,), dtype=torch.float32, device=x.device)
    rc = N.lib().tdl_depthwise_bwd(N.ptr(x), N.ptr(wf), N.ptr(dy), N.ptr(y), N.ptr(dx), N.ptr(dw), N.ptr(db), _i(n),
                                   _i(h), _i(w), _i(c), _i(p), _i(q), _i(r), _i(s), _i(stride), _i(dil), _i(pad_h),
                                   _i(pad_w), _i(int(relu)), N.stream_ptr())
    N.check(rc, "tdl_depthwise_bwd")
    return dx, dw, db


def upsample_fwd(x, out_hw, out: Optional[torch.Tensor] = None):
    """``out``: an [N, OH, OW, C] tensor to write into; it may be a channel slice of a wider buffer."""
    n, h, w, c = x.shape
    oh, ow = out_hw
    y = torch.empty((n, oh, ow, c), dtype=x.dtype, device=x.device) if out is None else out
    assert y.dtype == x.dtype and tuple(y.shape) == (n, oh, ow, c) and y.stride(3) == 1
    ld = y.stride(2)
    assert y.stride(1) == ow * ld and (n == 1 or y.stride(0) == oh * ow * ld), "out must be a channel slice"
    rc = N.lib().tdl_upsample_fwd_ld(N.ptr(x), N.ptr(y), _i(n), _i(h), _i(w), _i(c), _i(oh), _i(ow), _i64(ld),
                                     _i(int(x.dtype == torch.bfloat16)), N.stream_ptr())
    N.check(rc, "tdl_upsample_fwd_ld")
    return y


def upsample_bwd(dy, in_shape, out_hw):
    """``dy`` [N, OH, OW, C] may be a channel slice of a wider gradient buffer."""
    n, h, w, c = in_shape
    oh, ow = out_hw
    assert dy.stride(3) == 1
    ld = dy.stride(2)
    assert dy.stride(1) == ow * ld and (n == 1 or dy.stride(0) == oh * ow * ld), "dy must be a channel slice"
    dx = torch.zeros(in_shape, dtype=torch.float32, device=dy.device)
    rc = N.lib().tdl_upsample_bwd_ld(N.ptr(dy), N.ptr(dx), _i(n), _i(h), _i(w), _i(c), _i(oh), _i(ow), _i64(ld),
                                     _i(int(dy.dtype == torch.bfloat16)), N.stream_ptr())
    N.check(rc, "tdl_upsample_bwd_ld")
    return dx.to(dy.dtype)


def u8_to_bf16_normalize(x_u8: torch.Tensor, mean: torch.Tensor, inv_std: torch.Tensor, out: Optional[torch.Tensor] = None,
                         c_pad: Optional[int] = None) -> torch.Tensor:
    """uint8 NHWC -> (x/255 - mean) * inv_std as bf16 NHWC, optionally zero-padding channels to ``c_pad``."""
    n, h, w, c = x_u8.shape
    cp = c_pad or c
    if out is None:
        out = torch.empty((n, h, w, cp), dtype=torch.bfloat16, device=x_u8.device)
    rc = N.lib().tdl_u8_to_bf16_normalize(N.ptr(x_u8), N.ptr(out), _i64(n * h * w), _i(c), _i(cp), N.ptr(mean),
                                          N.ptr(inv_std), N.stream_ptr())
    N.check(rc, "tdl_u8_to_bf16_normalize")
    return out


def depthwise3x3_fwd(x, w9c, bias, mask, stride, dil, pad_h, pad_w, out_hw, relu_in, relu_out, want_stats=False):
    """Fast 3x3 depthwise: x [N,H,W,C] bf16, w9c [9, C] fp32 (tap-major).  ``mask``: zero outputs where mask <= 0.
    ``want_stats``: also return per-CTA partial BatchNorm statistics of the output, ``(y, partials[nparts,2,C])``, or
    ``(y, None)`` when the shape is not served by the TMA kernel."""
    _chk_bf16_nhwc(x, "x")
    n, h, w, c = x.shape
    p, q = out_hw
    y = torch.empty((n, p, q, c), dtype=torch.bfloat16, device=x.device)
    tma_ok = c % 8 == 0 and ((stride in (1, 2) and dil == 1) or (stride == 1 and dil == 2))
    if want_stats and tma_ok and mask is None:
        partials = torch.empty((2 * int(N.lib().tdl_num_sms()), 2, c), dtype=torch.float32, device=x.device)
        nparts = ctypes.c_int(0)
        rc = N.lib().tdl_depthwise3x3_tma_fwd_stats(N.ptr(x), N.ptr(w9c), N.ptr(bias), N.ptr(y), _i(n), _i(h), _i(w),
                                                    _i(c), _i(p), _i(q), _i(stride), _i(dil), _i(pad_h), _i(pad_w),
                                                    _i(int(relu_in)), _i(int(relu_out)), N.ptr(partials),
                                                    ctypes.byref(nparts), N.stream_ptr())
        N.check(rc, "tdl_depthwise3x3_tma_fwd_stats")
        return y, partials[:int(nparts.value)]
    fn = N.lib().tdl_depthwise3x3_tma_fwd if tma_ok else N.lib().tdl_depthwise3x3_fwd
    rc = fn(N.ptr(x), N.ptr(w9c), N.ptr(bias), N.ptr(mask), N.ptr(y), _i(n), _i(h), _i(w), _i(c), _i(p), _i(q),
            _i(stride), _i(dil), _i(pad_h), _i(pad_w), _i(int(relu_in)), _i(int(relu_out)), N.stream_ptr())
    N.check(rc, "tdl_depthwise3x3_tma_fwd" if tma_ok else "tdl_depthwise3x3_fwd")
    return (y, None) if want_stats else y


def depthwise3x3_dgrad_s1(dy, w9c, mask, dil, pad_h, pad_w, addend=None):
    """Data gradient of a stride-1 3x3 depthwise layer: ``w9c`` are the layer's own [9, C] fp32 weights (the kernel
    reads the taps in reverse order); ``pad_h/pad_w`` are the gradient problem's pads (2*dil - forward pad);
    ``mask``: zero where mask <= 0; ``addend`` (dx-shaped bf16): added after the mask inside the kernel.  Falls back
    to the generic kernel on a flipped copy for other dilations."""
    _chk_bf16_nhwc(dy, "dy")
    n, h, w, c = dy.shape
    if not (c % 8 == 0 and dil in (1, 2)):
        dx = depthwise3x3_fwd(dy, w9c.view(3, 3, c).flip(0, 1).contiguous().view(9, c), None, mask, 1, dil, pad_h,
                              pad_w, (h, w), False, False)
        return dx if addend is None else dx + addend
    dx = torch.empty_like(dy)
    if addend is not None:
        _chk_bf16_nhwc(addend, "addend")
        assert addend.shape == dy.shape
    rc = N.lib().tdl_depthwise3x3_tma_dgrad(N.ptr(dy), N.ptr(w9c), N.ptr(mask), N.ptr(addend), N.ptr(dx), _i(n), _i(h),
                                            _i(w), _i(c), _i(dil), _i(pad_h), _i(pad_w), N.stream_ptr())
    N.check(rc, "tdl_depthwise3x3_tma_dgrad")
    return dx


def depthwise3x3_wgrad(x, dy, ymask, stride, dil, pad_h, pad_w, relu_in, want_bias, dw_out=None, db_out=None):
    """dw [9, C] (and db [C]) of a 3x3 depthwise convolution.  ``dw_out`` / ``db_out`` (fp32, contiguous): the
    gradients are ADDED into these buffers instead of freshly zeroed ones (direct accumulation into ``.grad``)."""
    n, h, w, c = x.shape
    _, p, q, _ = dy.shape
    if dw_out is not None:
        assert dw_out.dtype == torch.float32 and dw_out.is_contiguous() and dw_out.numel() == 9 * c
        dw = dw_out.view(9, c)
    else:
        dw = torch.zeros((9, c), dtype=torch.float32, device=x.device)
    if not want_bias:
        db = None
    elif db_out is not None:
        assert db_out.dtype == torch.float32 and db_out.is_contiguous() and db_out.numel() == c
        db = db_out
    else:
        db = torch.zeros((c,), dtype=torch.float32, device=x.device)
    tma_ok = ymask is None and c % 8 == 0 and ((stride in (1, 2) and dil == 1) or (stride == 1 and dil == 2))
    if tma_ok:
        rc = N.lib().tdl_depthwise3x3_tma_wgrad(N.ptr(x), N.ptr(dy), N.ptr(dw), N.ptr(db), _i(n), _i(h), _i(w), _i(c),
                                                _i(p), _i(q), _i(stride), _i(dil), _i(pad_h), _i(pad_w),
                                                _i(int(relu_in)), N.stream_ptr())
        N.check(rc, "tdl_depthwise3x3_tma_wgrad")
        return dw, db
    rc = N.lib().tdl_depthwise3x3_wgrad(N.ptr(x), N.ptr(dy), N.ptr(ymask), N.ptr(dw), N.ptr(db), _i(n), _i(h), _i(w),
                                        _i(c), _i(p), _i(q), _i(stride), _i(dil), _i(pad_h), _i(pad_w),
                                        _i(int(relu_in)), N.stream_ptr())
    N.check(rc, "tdl_depthwise3x3_wgrad")
    return dw, db


def conv2d_smallk_fwd(x, w_krsc, bias, dil, pad_h, pad_w, out_hw):
    """K <= 4 output channels, stride 1 (csrc/smallk_conv.cu); ``w_krsc`` is the fp32 master weight."""
    _chk_bf16_nhwc(x, "x")
    n, h, w, c = x.shape
    K, R, S, _ = w_krsc.shape
    P, Q = out_hw
    y = torch.empty((n, P, Q, K), dtype=torch.bfloat16, device=x.device)
    rc = N.lib().tdl_conv2d_smallk_fwd(N.ptr(x), N.ptr(w_krsc), N.ptr(bias), N.ptr(y),
                                       _i(n), _i(h), _i(w), _i(c), _i(K), _i(R), _i(S), _i(dil), _i(pad_h), _i(pad_w),
                                       _i(P), _i(Q), N.stream_ptr())
    N.check(rc, "tdl_conv2d_smallk_fwd")
    return y


def conv2d_smallk_bwd_data(dy, w_krsc, in_shape, dil, pad_h, pad_w):
    n, h, w, c = in_shape
    K, R, S, _ = w_krsc.shape
    P, Q = dy.shape[1], dy.shape[2]
    dx = torch.empty(in_shape, dtype=torch.bfloat16, device=dy.device)
    rc = N.lib().tdl_conv2d_smallk_bwd_data(N.ptr(dy), N.ptr(w_krsc), N.ptr(dx), _i(n), _i(h), _i(w), _i(c), _i(K),
                                            _i(R), _i(S), _i(dil), _i(pad_h), _i(pad_w), _i(P), _i(Q), N.stream_ptr())
    N.check(rc, "tdl_conv2d_smallk_bwd_data")
    return dx


def conv2d_smallk_bwd_weight(x, dy, dw, dbias, dil, pad_h, pad_w):
    """Accumulates into ``dw`` (fp32 [K,R,S,C]) and, when given, ``dbias`` (fp32 [K])."""
    n, h, w, c = x.shape
    K, R, S, _ = dw.shape
    P, Q = dy.shape[1], dy.shape[2]
    rc = N.lib().tdl_conv2d_smallk_bwd_weight(N.ptr(x), N.ptr(dy), N.ptr(dw), N.ptr(dbias),
                                              _i(n), _i(h), _i(w), _i(c), _i(K), _i(R), _i(S), _i(dil), _i(pad_h),
                                              _i(pad_w), _i(P), _i(Q), N.stream_ptr())
    N.check(rc, "tdl_conv2d_smallk_bwd_weight")


def build_weight_prep_table(entries):
    """entries: list of (w fp32 [K,taps*C...], wf bf16, wd bf16 or None, K, taps, C).  Returns (descs, block_start,
    total_blocks) device tensors for :func:`conv2d_weight_prep_batched` (48-byte records, see include/tdl_b200.h)."""
    import numpy as np

    rec = np.zeros(len(entries), dtype=np.dtype([("w", "<u8"), ("wf", "<u8"), ("wd", "<u8"), ("K", "<i4"),
                                                 ("taps", "<i4"), ("C", "<i4"), ("ct", "<i4"), ("kt", "<i4"),
                                                 ("pad", "<i4")]))
    assert rec.dtype.itemsize == 48
    starts, total = [], 0
    dev = entries[0][0].device
    for i, (w, wf, wd, K, taps, C) in enumerate(entries):
        assert w.dtype == torch.float32 and w.is_contiguous() and wf.dtype == torch.bfloat16
        assert C % 2 == 0 and K % 2 == 0 and w.data_ptr() % 8 == 0 and wf.data_ptr() % 4 == 0
        ct, kt = (C + 63) // 64, (K + 63) // 64  # 64 x 64 tiles (csrc/conv_tcgen05.cu, weight_prep_batched_kernel)
        rec[i] = (w.data_ptr(), wf.data_ptr(), 0 if wd is None else wd.data_ptr(), K, taps, C, ct, kt, 0)
        starts.append(total)
        total += ct * kt * taps
    descs = torch.from_numpy(rec.view(np.uint8).copy()).to(dev)
    block_start = torch.tensor(starts, dtype=torch.int32).to(dev)
    return descs, block_start, total


def conv2d_weight_prep_batched(descs: torch.Tensor, block_start: torch.Tensor, total_blocks: int) -> None:
    rc = N.lib().tdl_conv2d_weight_prep_batched(N.ptr(descs), N.ptr(block_start), _i(block_start.numel()),
                                                _i(total_blocks), N.stream_ptr())
    N.check(rc, "tdl_conv2d_weight_prep_batched")


def conv_stem_pack(x: torch.Tensor) -> torch.Tensor:
    """[N,H,W,C<=4] bf16 -> space-to-depth planes [N, ceil(H/2), 2, ceil(W/2), 8] bf16 for the first-layer kernel."""
    _chk_bf16_nhwc(x, "x")
    n, h, w, c = x.shape
    planes = torch.empty((n, (h + 1) // 2, 2, (w + 1) // 2, 8), dtype=torch.bfloat16, device=x.device)
    rc = N.lib().tdl_conv_stem_pack(N.ptr(x), N.ptr(planes), _i(n), _i(h), _i(w), _i(c), N.stream_ptr())
    N.check(rc, "tdl_conv_stem_pack")
    return planes


def conv_stem_weight_prep(w_krsc: torch.Tensor, pad_h: int, pad_w: int) -> torch.Tensor:
    """fp32 master [K,R,S,C] -> bf16 [K, taps*16] for the space-to-depth problem."""
    K, R, S, C = w_krsc.shape
    elems = int(N.lib().tdl_conv_stem_weight_elems(_i(K), _i(R), _i(S), _i(pad_h), _i(pad_w)))
    out = torch.empty((K, elems // K), dtype=torch.bfloat16, device=w_krsc.device)
    rc = N.lib().tdl_conv_stem_weight_prep(N.ptr(w_krsc), N.ptr(out), _i(K), _i(R), _i(S), _i(C), _i(pad_h), _i(pad_w),
                                           N.stream_ptr())
    N.check(rc, "tdl_conv_stem_weight_prep")
    return out


def conv_stem_fprop(planes: torch.Tensor, w_packed: torch.Tensor, image_hw, K: int, R: int, S: int, pad_h: int,
                    pad_w: int, out_hw, want_stats: bool = False):
    """Stride-2 first-layer convolution on the tensor cores.  Returns y or (y, partials, nparts)."""
    n = planes.shape[0]
    h, w = image_hw
    p, q = out_hw
    y = torch.empty((n, p, q, K), dtype=torch.bfloat16, device=planes.device)
    if want_stats:
        partials = torch.empty((int(N.lib().tdl_num_sms()), 2, K), dtype=torch.float32, device=planes.device)
        nparts = ctypes.c_int(0)
        rc = N.lib().tdl_conv_stem_fprop(N.ptr(planes), N.ptr(w_packed), N.ptr(y), _i(n), _i(h), _i(w), _i(K), _i(R),
                                         _i(S), _i(pad_h), _i(pad_w), _i(p), _i(q), N.ptr(partials),
                                         ctypes.byref(nparts), N.stream_ptr())
        N.check(rc, "tdl_conv_stem_fprop")
        return y, partials, int(nparts.value)
    rc = N.lib().tdl_conv_stem_fprop(N.ptr(planes), N.ptr(w_packed), N.ptr(y), _i(n), _i(h), _i(w), _i(K), _i(R), _i(S),
                                     _i(pad_h), _i(pad_w), _i(p), _i(q), N.ptr(None), N.ptr(None), N.stream_ptr())
    N.check(rc, "tdl_conv_stem_fprop")
    return y


def conv_stem_wgrad(planes: torch.Tensor, dy: torch.Tensor, image_hw, C: int, R: int, S: int, pad_h: int, pad_w: int,
                    dw: Optional[torch.Tensor] = None, variant: int = 1) -> torch.Tensor:
    """First-layer weight gradient on the tensor cores from the packed image.  Returns fp32 [K,R,S,C]; when ``dw``
    is given the gradient is ADDED to it."""
    _chk_bf16_nhwc(dy, "dy")
    n, p, q, K = dy.shape
    h, w = image_hw
    elems = int(N.lib().tdl_conv_stem_weight_elems(_i(K), _i(R), _i(S), _i(pad_h), _i(pad_w)))
    scratch = torch.empty((elems,), dtype=torch.float32, device=dy.device)
    acc = dw is not None
    if dw is None:
        dw = torch.empty((K, R, S, C), dtype=torch.float32, device=dy.device)
    rc = N.lib().tdl_conv_stem_wgrad(N.ptr(planes), N.ptr(dy), N.ptr(scratch), N.ptr(dw), _i(n), _i(h), _i(w), _i(C),
                                     _i(K), _i(R), _i(S), _i(pad_h), _i(pad_w), _i(p), _i(q), _i(1 if acc else 0),
                                     _i(variant), N.stream_ptr())
    N.check(rc, "tdl_conv_stem_wgrad")
    return dw


def depthwise3x3_dgrad_s2(dy, w9c, xmask, in_shape, pad_h, pad_w):
    """Data gradient of a stride-2 3x3 depthwise conv; ``xmask`` (the forward input of a pre-activation ReLU) masks it."""
    n, h, w, c = in_shape
    _, p, q, _ = dy.shape
    dx = torch.empty(in_shape, dtype=torch.bfloat16, device=dy.device)
    rc = N.lib().tdl_depthwise3x3_dgrad_s2(N.ptr(dy), N.ptr(w9c), N.ptr(xmask), N.ptr(dx), _i(n), _i(h), _i(w), _i(c),
                                           _i(p), _i(q), _i(pad_h), _i(pad_w), N.stream_ptr())
    N.check(rc, "tdl_depthwise3x3_dgrad_s2")
    return dx


def conv1x1_fold_affine(w_kc: torch.Tensor, scale: torch.Tensor, shift: torch.Tensor,
                        bias: Optional[torch.Tensor] = None):
    """BatchNorm (no activation) folded into the 1x1 convolution that follows: returns (W * scale as bf16 [K, 1, C],
    bias + W @ shift as fp32 [K]).  ``w_kc``: fp32 master weights [K, C]."""
    K, C = w_kc.shape
    assert w_kc.dtype == torch.float32 and w_kc.is_contiguous()
    wf = torch.empty((K, 1, C), dtype=torch.bfloat16, device=w_kc.device)
    b = torch.empty((K,), dtype=torch.float32, device=w_kc.device)
    rc = N.lib().tdl_conv1x1_fold_affine(N.ptr(w_kc), N.ptr(scale), N.ptr(shift), N.ptr(bias), N.ptr(wf), N.ptr(b),
                                         _i(K), _i(C), N.stream_ptr())
    N.check(rc, "tdl_conv1x1_fold_affine")
    return wf, b


def conv1x1_fold_wgrad_fixup_(dw_kc: torch.Tensor, scale: torch.Tensor, shift: torch.Tensor, dzsum: torch.Tensor):
    """In place: dw[k][c] = scale[c] * dw[k][c] + shift[c] * dzsum[k] (weight gradient of the folded convolution);
    ``dzsum=None``: the column sums are known to be zero."""
    K, C = dw_kc.shape
    assert dw_kc.dtype == torch.float32 and dw_kc.is_contiguous()
    rc = N.lib().tdl_conv1x1_fold_wgrad_fixup(N.ptr(dw_kc), N.ptr(scale), N.ptr(shift), N.ptr(dzsum), _i(K), _i(C),
                                              N.stream_ptr())
    N.check(rc, "tdl_conv1x1_fold_wgrad_fixup")
    return dw_kc


def colsum_(dy2d: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
    """out[C] (fp32) += column sums of dy2d [rows, C] (bf16): a convolution's bias gradient in one pass."""
    rows, C = dy2d.shape
    rc = N.lib().tdl_colsum_bf16(N.ptr(dy2d), N.ptr(out), _i64(rows), _i(C), N.stream_ptr())
    N.check(rc, "tdl_colsum_bf16")
    return out
