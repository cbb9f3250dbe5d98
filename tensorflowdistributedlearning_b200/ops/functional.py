"""Autograd functions that bind the native sm_100a kernels, with the torch oracle as CPU path.

Dispatch rule (one rule, no backend matrix): a CUDA bf16 NHWC tensor whose channel counts meet
the kernels' 16-byte alignment goes to the native library; everything else (CPU towers used by
the plumbing tests, fp32 debugging, odd channel counts such as the 3-channel stem) runs the
plain-torch restatement in ``ops.reference``.  On a GPU the native library is mandatory --
``_native.lib()`` raises if it cannot be loaded, there is no silent fallback.
"""
from __future__ import annotations

import os
from typing import Optional, Tuple

import torch
import torch.nn.functional as F

from . import reference as R

# A data gradient that already carries the ReLU backward of the tensor it belongs to (applied in the conv epilogue):
# data_ptr of the gradient -> data_ptr of the 1-bit mask that was applied.  Consumed by that tensor's BatchNorm backward,
# which then skips its own masking and returns the gradient itself for the residual input.
_PREMASKED = {}
# A gradient whose tensor is the output of a convolution with a fused output ReLU and that ALREADY carries that ReLU's
# backward: data_ptr of the gradient -> (version, data_ptr of the activation).  Left by the BatchNorm that consumed the
# activation (its backward kernel zeroes the gradient where the activation is 0); the convolution then skips its own
# masking pass.
_RELU_DONE = {}
_LAST_AFFINE = None
_LAST_MEAN_OFFSET = None
# BatchNorm without activation in front of a 1x1 convolution is folded into that convolution's weights (TDL_DEFER_BN=0
# restores the separate normalise pass)
_DEFER_BN = os.environ.get("TDL_DEFER_BN", "1") != "0"
_LAST_RELU_MASK = None
# BatchNorm -> consumer convolution hand-off for the fused backward reductions: the BatchNorm's output tensor carries
# ``_tdl_bn_src`` (its input, batch statistics, affine, mask); the consumer's data-gradient kernel reduces
# (sum dz, sum dz * xhat) in its epilogue and leaves them here, keyed by the gradient's address and guarded by the
# tensor's version counter (autograd accumulating another consumer's gradient in place bumps it) and by the
# BatchNorm input's address.
_BN_PARTIALS = {}
_LAST_BN_SRC = None
_FUSE_BN_BWD = os.environ.get("TDL_FUSE_BN_BWD", "0") == "1"  # measured slower on B200: off by default


_FUSED_BN_HITS = 0
# BatchNorm with the statistics finalize folded into the apply kernels (cluster + DSMEM, csrc/bn.cu): one launch less
# per BatchNorm in forward and one less in backward -- numerically identical, but MEASURED SLOWER on B200 (the fold is
# a serial latency chain in front of the streaming loop and cluster launches stream slower: ResNet-50 19.3 -> 32.1
# ms/step, profiles/bn_fold_experiment.txt), so it is off unless TDL_BN_FOLD=1 asks for it.
_BN_FOLD = os.environ.get("TDL_BN_FOLD", "0") == "1"


def enable_defer_bn(flag: bool) -> None:
    """BatchNorm without activation in front of a 1x1 convolution folded into the convolution (default on)."""
    global _DEFER_BN
    _DEFER_BN = bool(flag)


def enable_bn_fold(flag: bool) -> None:
    global _BN_FOLD
    _BN_FOLD = bool(flag)


# Batched weight preparation (trainer-owned steps only): the bf16 fprop / dgrad copies of every convolution weight
# live in persistent buffers that ONE kernel refreshes at the start of a step (`begin_step`), instead of one small
# launch per layer inside the forward pass.  Prepared copies are valid only within the step that made them: the
# trainer calls `invalidate_prepared_weights()` after every step (eager or replayed), so an evaluation forward after
# an optimizer update prepares its own.
class _WeightPrepRegistry:
    """Persistent bf16 copies + the device table of one trainer.  The trainer holds a reference for as long as its
    captured graph may replay: the graph's nodes point at these buffers, so they must outlive any registry switch
    (a second trainer in the process) -- superseded tables are kept too (a few hundred bytes each)."""

    def __init__(self):
        self.entries = {}
        self.table = None
        self.retired = []
        self.zero_dx = {}  # pre-zeroed data-gradient buffers of strided convolutions with tap-less parity classes

    def drop_table(self):
        if self.table is not None:
            self.retired.append(self.table)
        self.table = None


_WPREP_ON = False
_WPREP = _WeightPrepRegistry()
_WPREP_STEP = 0


def enable_batched_weight_prep(flag: bool):
    """Starts a fresh registry (entries belong to the trainer that enables it) and returns it -- keep the reference
    while a CUDA graph captured under it may still be replayed."""
    global _WPREP_ON, _WPREP
    _WPREP_ON = bool(flag)
    _WPREP = _WeightPrepRegistry()
    return _WPREP


def invalidate_prepared_weights() -> None:
    global _WPREP_STEP
    _WPREP_STEP += 1


def _prepare_all_weights() -> None:
    from . import native_ops as O

    reg = _WPREP
    if not reg.entries:
        return
    if reg.table is None:
        if torch.cuda.is_current_stream_capturing():
            return  # a layer first seen during capture: its own launch serves it, the table waits for an eager step
        ents = list(reg.entries.values())
        reg.table = (O.build_weight_prep_table([(e["w"].detach(), e["wf"], e["wd"], e["K"], e["taps"], e["C"])
                                                for e in ents]), ents)
    (descs, starts, total), ents = reg.table
    O.conv2d_weight_prep_batched(descs, starts, total)
    _count(1)
    for e in ents:
        e["step"] = _WPREP_STEP


def enable_fused_bn_backward(flag: bool) -> None:
    global _FUSE_BN_BWD
    _FUSE_BN_BWD = bool(flag)


def fused_bn_backward_hits() -> int:
    """How many BatchNorm backward passes took their reductions from the producing data-gradient kernel."""
    return _FUSED_BN_HITS


def begin_step() -> None:
    """Drop per-step bookkeeping (call at the start of every training step)."""
    _PREMASKED.clear()
    _RELU_DONE.clear()
    _BN_PARTIALS.clear()
    if _WPREP_ON:
        invalidate_prepared_weights()
        _prepare_all_weights()


_STEM_NATIVE = True       # first-layer tcgen05 kernel (off = cuDNN for the C<=4 stem, used for A/B timing)
_FORCE_REFERENCE = False  # flipped by tests / the proxy baseline to route everything through torch ops
_LAUNCHES = 0             # native kernel launches issued through this module (bench "gpu_launches")


# Parameters whose gradient the native backward kernels accumulate straight into ``param.grad`` (a view of the flat
# gradient bucket) instead of returning it to autograd.  The distributed optimizer registers a callback per parameter
# so bucket readiness is still signalled (autograd's post-accumulate hooks do not fire for these).
_DIRECT_GRAD = False
_GRAD_READY = {}


def enable_direct_grad(flag: bool) -> None:
    """When on, conv / BatchNorm backward write fp32 gradients directly into existing ``.grad`` buffers."""
    global _DIRECT_GRAD
    _DIRECT_GRAD = bool(flag)


def register_grad_ready(param, fn) -> None:
    _GRAD_READY[id(param)] = fn


def unregister_grad_ready(param) -> None:
    _GRAD_READY.pop(id(param), None)


def _notify(param) -> None:
    fn = _GRAD_READY.get(id(param))
    if fn is not None:
        fn(param)


# Weight-gradient kernels only feed the optimizer step, so they can leave the critical path: with async wgrad on,
# they are launched on a side stream (forked after dy is ready) and overlap the memory-bound BatchNorm-backward /
# data-gradient chain that continues on the main stream.  `join_side_streams()` re-joins before the optimizer.
_ASYNC_WGRAD = False
_SIDE = {}
_SIDE_DIRTY = False


def enable_async_wgrad(flag: bool) -> None:
    global _ASYNC_WGRAD
    _ASYNC_WGRAD = bool(flag)


def _side_stream():
    dev = torch.cuda.current_device()
    st = _SIDE.get(dev)
    if st is None:
        st = _SIDE[dev] = torch.cuda.Stream(device=dev)
    return st


def join_side_streams(waiter=None) -> None:
    """Make ``waiter`` (default: the current stream) wait for every wgrad launched on the side stream."""
    global _SIDE_DIRTY
    if not _SIDE:
        return
    w = waiter or torch.cuda.current_stream()
    st = _SIDE.get(torch.cuda.current_device())
    if st is not None:
        w.wait_stream(st)
    if waiter is None:
        _SIDE_DIRTY = False


def _direct_ok(p) -> bool:
    return (_DIRECT_GRAD and p is not None and p.grad is not None and p.grad.dtype == torch.float32
            and p.grad.is_contiguous())


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


def _fold_wgrad_fixup(affine, dw_kc, dy, K):
    """Weight gradient of a convolution whose input BatchNorm was folded into it (see _ConvNative.forward): the kernel
    ran on the un-normalised input, dW = scale * dW_raw + shift x colsum(dy).  Same stream as the wgrad before it."""
    if affine is None:
        return
    from . import native_ops as O

    if affine[2]:
        # the convolution feeds a training-mode BatchNorm: dy is that BatchNorm's input gradient, whose column sums are
        # zero by construction -- the shift term vanishes and no pass over dy is needed
        O.conv1x1_fold_wgrad_fixup_(dw_kc, affine[0], affine[1], None)
        _count(1)
        return
    dzsum = O.colsum_(dy.view(-1, K), torch.zeros(K, dtype=torch.float32, device=dy.device))
    O.conv1x1_fold_wgrad_fixup_(dw_kc, affine[0], affine[1], dzsum)
    _count(2)


class _ConvNative(torch.autograd.Function):
    """tcgen05 implicit-GEMM forward; dgrad = same kernel on flipped/transposed weights; wgrad = MN-major kernel."""

    @staticmethod
    def forward(ctx, x, weight, bias, stride, dil, pads, out_hw, want_stats=False, want_tap=False, residual=None,
                relu_out=False):
        from . import native_ops as O

        ctx.weight_param, ctx.bias_param, ctx.want_tap = weight, bias, want_tap
        ctx.relu_out, ctx.has_residual = bool(relu_out), residual is not None
        # the statistics output is not differentiable and the tap's gradient may be absent: do not let autograd
        # launch a zero-fill for each of them every step
        ctx.set_materialize_grads(False)
        ctx.y_shape = (x.shape[0], out_hw[0], out_hw[1], weight.shape[0])
        # the 1-bit ReLU mask of the input tensor, when its producer left one (BatchNorm + residual + ReLU)
        ctx.in_mask = getattr(x, "_tdl_relu_mask", None) if want_tap else None
        ctx.bn_src = getattr(x, "_tdl_bn_src", None) if _FUSE_BN_BWD else None

        K, Rk, Sk, C = weight.shape
        taps = Rk * Sk
        need_dgrad = ctx.needs_input_grad[0]
        need_wd = need_dgrad  # stride 1: the whole flipped copy; stride 2: each parity class walks a subset of it
        ent = _WPREP.entries.get(id(weight)) if _WPREP_ON else None
        if (ent is not None and ent["w"] is weight and ent["step"] == _WPREP_STEP
                and (ent["wd"] is not None or not need_wd)):
            wf, wd = ent["wf"], (ent["wd"] if need_wd else None)  # refreshed by this step's one batched launch
            prep_launches = 0
        else:
            wf = torch.empty((K, taps, C), dtype=torch.bfloat16, device=x.device)
            wd = torch.empty((C, taps, K), dtype=torch.bfloat16, device=x.device) if need_wd else None
            O.conv2d_weight_prep(weight.detach(), wf, wd, K, taps, C)
            prep_launches = 1
            stale = ent is not None and (ent["w"] is not weight or (need_wd and ent["wd"] is None))
            if (_WPREP_ON and (ent is None or stale) and isinstance(weight, torch.nn.Parameter)
                    and weight.is_contiguous() and not torch.cuda.is_current_stream_capturing()):
                # (a replaced entry's buffers stay alive through the retired table that still lists them)
                _WPREP.entries[id(weight)] = dict(w=weight, wf=wf, wd=wd, K=K, taps=taps, C=C, step=_WPREP_STEP)
                _WPREP.drop_table()
        pt, pb, pl, pr = pads
        partials = None
        # input = a BatchNorm output that was never written (batch_norm(defer_apply=True)): fold its scale / shift
        # into this 1x1 convolution's weights; backward corrects the weight gradient (csrc/bn.cu, fold_* kernels)
        ctx.affine = getattr(x, "_tdl_affine", None)
        bias_eff = None if bias is None else bias.detach()
        if ctx.affine is not None:
            if not (taps == 1 and stride == 1 and pads == (0, 0, 0, 0) and residual is None and not want_tap):
                raise RuntimeError("a deferred BatchNorm output may only feed a plain 1x1 stride-1 convolution")
            wf, bias_eff = O.conv1x1_fold_affine(weight.detach().view(K, C), ctx.affine[0], ctx.affine[1], bias_eff)
            prep_launches += 1
            if ctx.affine[2]:
                # the output goes into a training-mode BatchNorm, which removes any per-channel constant: leave the
                # folded bias out of the tensor (the bias epilogue costs 45 % on these layers) and hand it to that
                # BatchNorm's moving-mean update instead (conv2d() tags the output)
                global _LAST_MEAN_OFFSET
                _LAST_MEAN_OFFSET = bias_eff
                bias_eff = None
        if want_stats:
            y, partials, nparts = O.conv2d_fprop_stats(x, wf, R=Rk, S=Sk, stride=stride, dil=dil, pad_h=pt, pad_w=pl,
                                                       out_hw=out_hw, bias=bias_eff,
                                                       residual=(None if residual is None
                                                                 else residual.detach().contiguous()), relu=relu_out)
            partials = partials[:nparts]
        else:
            # (+ residual) (+ ReLU) in the epilogue: the reference unit's relu(shortcut + conv3(...)), core/resnet.py:148
            y = O.conv2d_fprop(x, wf, R=Rk, S=Sk, stride=stride, dil=dil, pad_h=pt, pad_w=pl, out_hw=out_hw,
                               bias=bias_eff,
                               residual=None if residual is None else residual.detach().contiguous(), relu=relu_out)
        _count(1 + prep_launches)
        ctx.save_for_backward(x, wd if wd is not None else wf, y if relu_out else None)
        ctx.cfg = (K, Rk, Sk, C, stride, dil, pads, tuple(x.shape), bias is not None, wd is not None)
        outs = [y]
        if want_stats:
            ctx.mark_non_differentiable(partials)
            outs.append(partials)
        if want_tap:
            # an alias of the input for an identity shortcut: its gradient comes back to THIS node, where the
            # data-gradient kernel adds it in its epilogue instead of autograd launching a separate add
            outs.append(x.view_as(x))
        ctx.n_out = len(outs)
        ctx.has_stats = want_stats
        return outs[0] if len(outs) == 1 else tuple(outs)

    @staticmethod
    def backward(ctx, dy, *rest):
        from . import native_ops as O

        global _SIDE_DIRTY
        x, wsaved, y_relu = ctx.saved_tensors
        K, Rk, Sk, C, stride, dil, pads, xshape, has_bias, has_wd = ctx.cfg
        pt, pb, pl, pr = pads
        n, h, w, _ = xshape
        if dy is None:  # only the tap / statistics outputs were used downstream
            dy = torch.zeros(ctx.y_shape, dtype=x.dtype, device=x.device)
        dy = dy.contiguous()
        if ctx.relu_out:  # backward of the fused output ReLU: everything below, and the residual branch, sees dy * (y > 0)
            done = _RELU_DONE.pop(dy.data_ptr(), None)
            if done is None or done != (dy._version, y_relu.data_ptr()):
                dy = torch.ops.aten.threshold_backward(dy, y_relu, 0)
            # else: the only consumer of y was a BatchNorm whose backward kernel already zeroed dy where y == 0
        d_tap = rest[-1] if (ctx.want_tap and rest and rest[-1] is not None) else None
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            if has_wd and stride == 1:
                mask = ctx.in_mask if (d_tap is not None and C % 32 == 0) else None
                src = ctx.bn_src
                geo = dict(R=Rk, S=Sk, dil=dil, pad_h=(Rk - 1) * dil - pt, pad_w=(Sk - 1) * dil - pl, out_hw=(h, w))
                res = None if d_tap is None else d_tap.contiguous()
                partials = None
                if src is not None and src["x"].numel() == n * h * w * C and (ctx.want_tap == (mask is not None)):
                    # the tensor this gradient belongs to came straight out of a BatchNorm: reduce that BatchNorm's
                    # backward sums here, in the epilogue (tap path: mask applied; plain chain: mask recomputed)
                    rc = src["relu"] and not src["has_res"] and mask is None
                    if mask is not None or rc or not src["relu"]:
                        dx, partials = O.conv2d_dgrad_bnstats(dy, wsaved, bn_x=src["x"], bn_mean=src["mean"],
                                                              bn_invstd=src["invstd"], residual=res, out_mask=mask,
                                                              bn_scale=src["scale"] if rc else None,
                                                              bn_shift=src["shift"] if rc else None, **geo)
                        _BN_PARTIALS[dx.data_ptr()] = (dx._version, src["x"].data_ptr(), partials, mask is not None)
                if partials is None:
                    dx = O.conv2d_fprop(dy, wsaved, stride=1, residual=res, out_mask=mask, **geo)
                if mask is not None:  # this IS the total gradient of x (tap), with x's ReLU backward already applied
                    # (version-guarded: autograd summing another consumer's gradient into dx in place keeps the
                    # address but bumps the version, and the sum is NOT masked as a whole any more)
                    _PREMASKED[dx.data_ptr()] = (dx._version, mask.data_ptr())
                d_tap = None
                _count(1)
            else:
                # strided data gradient: one tcgen05 launch per input-parity class, strided TMA stores
                dx = O.conv2d_dgrad_strided(dy, wsaved, (n, h, w, C), R=Rk, S=Sk, stride=stride, pad_h=pt, pad_w=pl,
                                            zero_buffers=_WPREP.zero_dx if _WPREP_ON else None,
                                            zero_key=(id(ctx.weight_param)
                                                      if isinstance(ctx.weight_param, torch.nn.Parameter) else None))
                _count(stride * stride)
        if d_tap is not None and dx is not None:
            dx = dx + d_tap
        if ctx.needs_input_grad[1] and C <= 16:
            # tiny Cin (channel-padded RGB stem): a 128x64 tensor-core tile would be >90% padding, cuDNN's
            # small-channel wgrad is the better tool for this one layer
            w4 = ctx.weight_param.detach().to(x.dtype).permute(0, 3, 1, 2)  # only its shape matters here
            if w4.shape[1] != C:
                w4 = F.pad(w4, (0, 0, 0, 0, 0, C - w4.shape[1]))
            xin = x.permute(0, 3, 1, 2)
            if pt != pb or pl != pr:
                xin = F.pad(xin, (pl, pr, pt, pb))
                padding = [0, 0]
            else:
                padding = [pt, pl]
            _, gw, _ = torch.ops.aten.convolution_backward(dy.permute(0, 3, 1, 2), xin, w4, None, [stride, stride],
                                                           padding, [dil, dil], False, [0, 0], 1, [False, True, False])
            dw = gw.permute(0, 2, 3, 1).float().contiguous()
        elif ctx.needs_input_grad[1]:
            wp = ctx.weight_param
            if _direct_ok(wp):
                if _ASYNC_WGRAD:
                    side = _side_stream()
                    side.wait_stream(torch.cuda.current_stream())  # dy (and x) are complete on the main stream
                    with torch.cuda.stream(side):
                        O.conv2d_wgrad(x, dy, wp.grad.view(K, Rk * Sk, C), R=Rk, S=Sk, stride=stride, dil=dil,
                                       pad_h=pt, pad_w=pl)
                        _fold_wgrad_fixup(ctx.affine, wp.grad.view(K, C) if ctx.affine is not None else None, dy, K)
                    x.record_stream(side)   # keep the operands alive until the side stream is done with them
                    dy.record_stream(side)
                    _SIDE_DIRTY = True
                else:
                    O.conv2d_wgrad(x, dy, wp.grad.view(K, Rk * Sk, C), R=Rk, S=Sk, stride=stride, dil=dil, pad_h=pt,
                                   pad_w=pl)
                    _fold_wgrad_fixup(ctx.affine, wp.grad.view(K, C) if ctx.affine is not None else None, dy, K)
                _notify(wp)
            else:
                dw = torch.zeros((K, Rk * Sk, C), dtype=torch.float32, device=x.device)
                O.conv2d_wgrad(x, dy, dw, R=Rk, S=Sk, stride=stride, dil=dil, pad_h=pt, pad_w=pl)
                _fold_wgrad_fixup(ctx.affine, dw.view(K, C) if ctx.affine is not None else None, dy, K)
                dw = dw.view(K, Rk, Sk, C)
            _count(1)
        if has_bias and ctx.needs_input_grad[2]:
            bp = ctx.bias_param
            if K % 8 == 0:  # one pass over dy (bf16 -> fp32 column sums) instead of a cast plus a generic reduction
                if _direct_ok(bp):
                    if _ASYNC_WGRAD:  # like the weight gradient, nothing downstream in backward reads it
                        side = _side_stream()
                        side.wait_stream(torch.cuda.current_stream())
                        with torch.cuda.stream(side):
                            O.colsum_(dy.view(-1, K), bp.grad)
                        dy.record_stream(side)
                        _SIDE_DIRTY = True
                    else:
                        O.colsum_(dy.view(-1, K), bp.grad)
                    _notify(bp)
                else:
                    db = O.colsum_(dy.view(-1, K), torch.zeros(K, dtype=torch.float32, device=dy.device))
                _count(1)
            else:
                db = dy.float().sum((0, 1, 2))
        return dx, dw, db, None, None, None, None, None, None, (dy if ctx.has_residual else None), None


class _ConvStem(torch.autograd.Function):
    """First layer: stride-2 conv on <= 4 input channels (7x7 / 3x3 roots, core/resnet.py:164,219-224;
    core/xception.py:352).  Forward = space-to-depth pack + tcgen05 kernel with the haloed patch resident in shared
    memory (csrc/conv_tcgen05.cu, kStem); BatchNorm partial statistics optionally fused.  The weight gradient runs on
    the same packed image (stem_wgrad_kernel: one output row per GEMM-K slab, column taps as overlapping operand
    groups); only a requested INPUT gradient (never the case for an image) falls back to the library."""

    @staticmethod
    def forward(ctx, x, weight, pads, out_hw, want_stats):
        from . import native_ops as O

        K, Rk, Sk, C = weight.shape
        pt, pb, pl, pr = pads
        planes = O.conv_stem_pack(x)
        wp = O.conv_stem_weight_prep(weight.detach().contiguous(), pt, pl)
        res = O.conv_stem_fprop(planes, wp, (x.shape[1], x.shape[2]), K, Rk, Sk, pt, pl, out_hw, want_stats)
        _count(3)
        ctx.save_for_backward(x, weight, planes)
        ctx.set_materialize_grads(False)  # no zero-fill for the (non-differentiable) statistics output
        ctx.weight_param = weight
        ctx.cfg = (pads, want_stats)
        if want_stats:
            y, partials, nparts = res
            partials = partials[:nparts]
            ctx.mark_non_differentiable(partials)
            return y, partials
        return res

    @staticmethod
    def backward(ctx, dy, *rest):
        x, weight, planes = ctx.saved_tensors
        if dy is None:
            return None, None, None, None, None
        (pt, pb, pl, pr), _ = ctx.cfg
        K, Rk, Sk, C = weight.shape
        need_dx, need_dw = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        if need_dw and not need_dx:
            from . import native_ops as O

            wp = ctx.weight_param
            dy = dy.contiguous()
            hw = (x.shape[1], x.shape[2])
            _count(3)
            if _direct_ok(wp):
                O.conv_stem_wgrad(planes, dy, hw, C, Rk, Sk, pt, pl, dw=wp.grad)
                _notify(wp)
                return None, None, None, None, None
            return None, O.conv_stem_wgrad(planes, dy, hw, C, Rk, Sk, pt, pl), None, None, None
        w4 = weight.detach().to(x.dtype).permute(0, 3, 1, 2)
        xin = x.permute(0, 3, 1, 2)
        if pt != pb or pl != pr:
            xin = F.pad(xin, (pl, pr, pt, pb))
            padding = [0, 0]
        else:
            padding = [pt, pl]
        gx, gw, _ = torch.ops.aten.convolution_backward(dy.contiguous().permute(0, 3, 1, 2), xin, w4, None, [2, 2], padding,
                                                        [1, 1], False, [0, 0], 1, [need_dx, need_dw, False])
        dx = dw = None
        if need_dx:
            gx = gx[:, :, pt:pt + x.shape[1], pl:pl + x.shape[2]] if padding == [0, 0] else gx
            dx = gx.permute(0, 2, 3, 1).contiguous()
        if need_dw:
            dw = gw.permute(0, 2, 3, 1).float().contiguous()
        return dx, dw, None, None, None


class _ConvSmallK(torch.autograd.Function):
    """1..4 output channels, stride 1: CUDA-core dot-product kernels (csrc/smallk_conv.cu; SURVEY K4 -- the
    decoder's 3x3 conv to one logit channel, core/resnet.py:487-492)."""

    @staticmethod
    def forward(ctx, x, weight, bias, dil, pads, out_hw):
        from . import native_ops as O

        ctx.weight_param, ctx.bias_param = weight, bias
        pt, pb, pl, pr = pads
        wd = weight.detach().contiguous()
        y = O.conv2d_smallk_fwd(x, wd, None if bias is None else bias.detach(), dil, pt, pl, out_hw)
        _count(1)
        ctx.save_for_backward(x, wd)
        ctx.cfg = (dil, pt, pl, tuple(x.shape))
        return y

    @staticmethod
    def backward(ctx, dy):
        from . import native_ops as O

        x, wd = ctx.saved_tensors
        dil, pt, pl, xshape = ctx.cfg
        dy = dy.contiguous()
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = O.conv2d_smallk_bwd_data(dy, wd, xshape, dil, pt, pl)
            _count(1)
        wp, bp = ctx.weight_param, ctx.bias_param
        want_w = ctx.needs_input_grad[1]
        want_b = bp is not None and ctx.needs_input_grad[2]
        if want_w:
            direct_w = _direct_ok(wp)
            direct_b = want_b and _direct_ok(bp)
            gw = wp.grad if direct_w else torch.zeros_like(wd)
            gb = None
            if want_b:
                gb = bp.grad if direct_b else torch.zeros(wd.shape[0], dtype=torch.float32, device=x.device)
            if _ASYNC_WGRAD and direct_w and (direct_b or not want_b):
                # first node of backward: its weight gradient overlaps everything that follows on the main stream
                global _SIDE_DIRTY
                side = _side_stream()
                side.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(side):
                    O.conv2d_smallk_bwd_weight(x, dy, gw, gb, dil, pt, pl)
                x.record_stream(side)
                dy.record_stream(side)
                _SIDE_DIRTY = True
            else:
                O.conv2d_smallk_bwd_weight(x, dy, gw, gb, dil, pt, pl)
            _count(1)
            if direct_w:
                _notify(wp)
            else:
                dw = gw
            if want_b:
                if direct_b:
                    _notify(bp)
                else:
                    db = gb
        elif want_b:
            db = dy.float().sum((0, 1, 2))
        return dx, dw, db, None, None, None


def conv2d(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None, stride: int = 1,
           dil: int = 1, padding="SAME", want_stats: bool = False, want_tap: bool = False,
           residual: Optional[torch.Tensor] = None, relu: bool = False):
    """NHWC convolution, weight [K, R, S, C] (fp32 master).  TF 'SAME' geometry by default.

    ``residual`` / ``relu``: ``relu?(conv(x) + bias + residual)`` in the convolution's own epilogue (not combinable
    with ``want_tap``; with ``want_stats`` the statistics are those of the stored sum, and need the residual).

    ``want_stats=True`` returns ``(y, stats)`` where ``stats`` is either None (torch path) or the per-CTA partial
    BatchNorm statistics the tcgen05 kernel accumulated in its epilogue; pass it to :func:`batch_norm`.
    ``want_tap=True`` returns ``(y, stats_or_None, x_alias)``: feed ``x_alias`` to an identity shortcut and the
    shortcut's gradient is added inside the data-gradient kernel's epilogue."""
    K, Rk, Sk, C = weight.shape
    n, h, w, c = x.shape
    if c > C:  # channel-padded input (e.g. RGB padded to 8 for the TMA path): zero-pad the weight to match
        weight = F.pad(weight, (0, c - C))
        C = c
    assert c == C, f"channel mismatch: input {c} vs weight {C}"
    P, Q, pt, pb, pl, pr = conv_pads(h, w, Rk, Sk, stride, dil, padding)
    if getattr(x, "_tdl_affine", None) is not None and not (_native_ok(x) and C % 8 == 0 and K % 8 == 0 and K > 4):
        raise RuntimeError("a deferred BatchNorm output reached a convolution that cannot fold it")
    if (_native_ok(x) and c <= 4 and stride == 2 and dil == 1 and bias is None and K % 8 == 0 and K <= 64
            and max(Rk, Sk) <= 7 and not want_tap and _STEM_NATIVE and residual is None and not relu):
        out = _ConvStem.apply(x.contiguous(), weight, (pt, pb, pl, pr), (P, Q), want_stats)
        return out
    if (_native_ok(x) and K <= 4 and stride == 1 and Rk * Sk <= 9 and C % 8 == 0 and C <= 2048
            and not (want_stats or want_tap) and residual is None and not relu):
        return _ConvSmallK.apply(x.contiguous(), weight, bias, dil, (pt, pb, pl, pr), (P, Q))
    if _native_ok(x) and C % 8 == 0 and K % 8 == 0 and stride <= 2:
        if want_stats and residual is not None and not want_tap:
            # statistics of relu?(conv + bias + residual) as stored: for the BatchNorm that consumes the sum next
            return _ConvNative.apply(x.contiguous(), weight, bias, stride, dil, (pt, pb, pl, pr), (P, Q), True, False,
                                     residual, relu)
        if (want_stats or want_tap) and residual is None and not relu:
            global _LAST_MEAN_OFFSET
            _LAST_MEAN_OFFSET = None
            out = _ConvNative.apply(x.contiguous(), weight, bias, stride, dil, (pt, pb, pl, pr), (P, Q), want_stats,
                                    want_tap)
            if _LAST_MEAN_OFFSET is not None:
                if not (want_stats and not want_tap):
                    raise RuntimeError("a folded bias left out of the output needs the statistics-producing form")
                out[0]._tdl_mean_offset = _LAST_MEAN_OFFSET
                _LAST_MEAN_OFFSET = None
            if want_stats and want_tap:
                return out
            if want_tap:
                return out[0], None, out[1]
            return out
        assert not want_tap, "want_tap is not combinable with residual / relu"
        if residual is not None or relu:
            y = _ConvNative.apply(x.contiguous(), weight, bias, stride, dil, (pt, pb, pl, pr), (P, Q), False, False,
                                  residual, relu)
            if relu:
                y._tdl_relu_out = True  # y >= 0, and the producing node owns the backward of that ReLU
        else:
            y = _ConvNative.apply(x.contiguous(), weight, bias, stride, dil, (pt, pb, pl, pr), (P, Q))
        return (y, None) if want_stats else y
    # torch path (cuDNN on GPU through a channels-last view; plain CPU conv on CPU towers)
    wq = weight.to(x.dtype)
    bq = None if bias is None else bias.to(x.dtype)
    xin = x.permute(0, 3, 1, 2)
    if pt == pb and pl == pr:
        y = F.conv2d(xin, wq.permute(0, 3, 1, 2), bq, stride=stride, padding=(pt, pl), dilation=dil)
    else:
        y = F.conv2d(F.pad(xin, (pl, pr, pt, pb)), wq.permute(0, 3, 1, 2), bq, stride=stride, dilation=dil)
    y = y.permute(0, 2, 3, 1)
    if residual is not None:
        y = y + residual
    if relu:
        y = F.relu(y)
    if want_tap:
        return y, None, x
    return (y, None) if want_stats else y


# ---------------------------------------------------------------------------------------------
# batch norm (+ residual) (+ ReLU)
# ---------------------------------------------------------------------------------------------
class _BNNative(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, running_mean, running_var, residual, decay, eps, relu, training, partials=None,
                want_tap=False, defer=False):
        from . import native_ops as O

        ctx.gamma_param, ctx.beta_param = gamma, beta
        ctx.want_tap = bool(want_tap)
        ctx.x_is_relu_out = bool(getattr(x, "_tdl_relu_out", False))
        ctx.set_materialize_grads(False)

        shape = x.shape
        C = shape[-1]
        x2 = x.reshape(-1, C)
        r2 = None if residual is None else residual.reshape(-1, C)
        need_z = training and relu and residual is not None
        if training and _BN_FOLD:
            if partials is None:
                partials, nparts = O.bn_partial_stats(x2)
                _count(1)
            else:
                nparts = partials.shape[0]
            z, zmask, mean, invstd, scale, shift = O.bn_apply_fused(x2, partials, nparts, x2.shape[0], gamma, beta,
                                                                    running_mean, running_var, decay, eps, r2, relu,
                                                                    want_mask=need_z)
            _count(1)
            ctx.save_for_backward(x2, zmask, scale, shift, mean, invstd)
            ctx.cfg = (relu, residual is not None, training, shape)
            global _LAST_RELU_MASK, _LAST_BN_SRC
            _LAST_RELU_MASK = zmask
            _LAST_BN_SRC = dict(x=x2, mean=mean, invstd=invstd, scale=scale, shift=shift, relu=bool(relu),
                                has_res=residual is not None)
            return (z.view(shape), x.view_as(x)) if want_tap else z.view(shape)
        mean_offset = getattr(x, "_tdl_mean_offset", None)
        if mean_offset is not None and not (training and partials is not None and not _BN_FOLD):
            raise RuntimeError("the producer left a folded bias out of this tensor: training-mode BatchNorm with the "
                               "producer's partial statistics is the only valid consumer")
        if training and partials is not None:
            mean, invstd, scale, shift = O.bn_finalize_partials(partials, partials.shape[0], x2.shape[0], gamma, beta,
                                                                running_mean, running_var, decay, eps, mean_offset)
            _count(1)
        elif training:
            mean, invstd, scale, shift = O.bn_fwd_stats(x2, gamma, beta, running_mean, running_var, decay, eps)
            _count(2)
        else:
            scale, shift = O.bn_infer_scale_shift(gamma, beta, running_mean, running_var, eps)
            mean = invstd = None
            _count(1)
        if defer:
            # The only consumer is a 1x1 convolution that folds scale / shift into its weights (_ConvNative): the
            # normalised tensor is never written.  The output is an alias of the input; batch_norm() tags it.
            assert training and not relu and residual is None and not want_tap
            z, zmask = x2, None
        elif need_z:
            # with a residual the ReLU mask cannot be recomputed from x alone: keep it as 1 bit per element (written by
            # the same kernel) instead of making both backward kernels re-read the whole activation
            z, zmask = O.bn_apply(x2, scale, shift, r2, relu, want_mask=True)
            _count(1)
        else:
            z, zmask = O.bn_apply(x2, scale, shift, r2, relu), None
            _count(1)
        if training:
            ctx.save_for_backward(x2, zmask, scale, shift, mean, invstd)
        ctx.cfg = (relu, residual is not None, training, shape)
        _LAST_RELU_MASK = zmask
        _LAST_BN_SRC = (dict(x=x2, mean=mean, invstd=invstd, scale=scale, shift=shift, relu=bool(relu),
                             has_res=residual is not None) if training else None)
        # want_tap: an alias of the input for an identity shortcut -- its gradient comes back to THIS node and is added
        # inside the backward apply kernel (no separate elementwise add; core/resnet.py:71-87,126-148)
        if defer:
            global _LAST_AFFINE
            _LAST_AFFINE = (scale, shift, defer == 2)
            return x.view_as(x)
        return (z.view(shape), x.view_as(x)) if want_tap else z.view(shape)

    @staticmethod
    def backward(ctx, dz, d_tap=None):
        from . import native_ops as O

        relu, has_res, training, shape = ctx.cfg
        if not training:
            raise RuntimeError("backward through inference-mode BatchNorm is not supported")
        x2, z2, scale, shift, mean, invstd = ctx.saved_tensors
        C = shape[-1]
        if dz is None:  # only the alias was used downstream
            dz = torch.zeros(shape, dtype=x2.dtype, device=x2.device)
        dz2 = dz.contiguous().view(-1, C)
        addend = None if d_tap is None else d_tap.contiguous().view(-1, C)
        # the tap makes this node's dx the gradient of the input from ALL of this unit's uses of it; when the input is
        # itself a ReLU output, that ReLU's backward is applied here as well (the producer is told through _RELU_DONE)
        mask_by_x = ctx.want_tap and ctx.x_is_relu_out and not _BN_FOLD
        g, b = ctx.gamma_param, ctx.beta_param
        direct = _direct_ok(g) and _direct_ok(b)
        pm = _PREMASKED.pop(dz2.data_ptr(), None)
        premasked = (z2 is not None and has_res and pm is not None and pm[0] == dz2._version
                     and pm[1] == z2.data_ptr())  # otherwise mask here: masking twice is harmless (idempotent)
        partials = None
        ent = _BN_PARTIALS.pop(dz2.data_ptr(), None)
        if ent is not None and ent[0] == dz2._version and ent[1] == x2.data_ptr() and ent[3] == premasked:
            partials = ent[2]  # the kernel that produced dz already reduced (sum dz, sum dz * xhat)
            global _FUSED_BN_HITS
            _FUSED_BN_HITS += 1
        bwd = O.bn_bwd_fused if _BN_FOLD else O.bn_bwd
        extra = {} if _BN_FOLD else dict(addend=addend, mask_by_x=mask_by_x)
        if premasked:
            # the consumer's data-gradient kernel already applied this layer's ReLU mask: no masking here, and the
            # residual branch's gradient is the incoming gradient itself (nothing to write)
            dx, _, dgamma, dbeta = bwd(dz2, x2, None, scale, shift, mean, invstd, False, False,
                                       acc_dgamma=g.grad if direct else None,
                                       acc_dbeta=b.grad if direct else None, partials=partials, **extra)
            dres = dz2
        elif has_res and not relu:
            # no activation after the sum: the residual branch's gradient is the incoming gradient, nothing to write
            dx, _, dgamma, dbeta = bwd(dz2, x2, None, scale, shift, mean, invstd, False, False,
                                       acc_dgamma=g.grad if direct else None,
                                       acc_dbeta=b.grad if direct else None, partials=partials, **extra)
            dres = dz2
        else:
            dx, dres, dgamma, dbeta = bwd(dz2, x2, z2, scale, shift, mean, invstd, relu, has_res,
                                          acc_dgamma=g.grad if direct else None,
                                          acc_dbeta=b.grad if direct else None, partials=partials, **extra)
        if _BN_FOLD and addend is not None:
            dx = dx + addend
        if mask_by_x:
            _RELU_DONE[dx.data_ptr()] = (dx._version, x2.data_ptr())
        _count((3 if partials is not None else 4) - (1 if _BN_FOLD else 0))
        if direct:
            _notify(g)
            _notify(b)
            dgamma = dbeta = None
        return (dx.view(shape), dgamma if ctx.needs_input_grad[1] else None,
                dbeta if ctx.needs_input_grad[2] else None, None, None,
                dres.view(shape) if has_res else None, None, None, None, None, None, None, None)


def batch_norm(x: torch.Tensor, gamma, beta, running_mean, running_var, *, decay: float, eps: float,
               training: bool, relu: bool = False, residual: Optional[torch.Tensor] = None,
               stats: Optional[torch.Tensor] = None, want_tap: bool = False, defer_apply: bool = False):
    """y = relu?(BN(x) + residual?) over the last (channel) axis with TF fused-BN semantics.
    ``defer_apply=True`` (training, no activation, no residual; the ONLY consumer must be a 1x1 stride-1 convolution
    through :func:`conv2d`): the normalise pass is skipped, the result is an alias of ``x`` tagged with (scale, shift)
    and the convolution folds them into its weights (Xception depthwise -> BN -> pointwise).  ``defer_apply=2``
    additionally promises that the convolution's OUTPUT feeds a training-mode BatchNorm (so its output gradient has
    zero column sums and the weight-gradient correction needs no pass over it).
    ``stats``: partial statistics of ``x`` from ``conv2d(..., want_stats=True)`` (skips the statistics pass).
    ``want_tap=True`` -> (y, x_alias): feed ``x_alias`` to an identity shortcut and its gradient is added inside this
    BatchNorm's backward kernel instead of by a separate elementwise add."""
    C = x.shape[-1]
    if _native_ok(x) and C % 8 == 0 and gamma is not None and beta is not None:
        res = None if residual is None else residual.contiguous()
        global _LAST_RELU_MASK, _LAST_BN_SRC
        _LAST_RELU_MASK = _LAST_BN_SRC = None
        tap = None
        xin = x.contiguous()
        if xin is not x and getattr(x, "_tdl_relu_out", False):
            xin._tdl_relu_out = True
        global _LAST_AFFINE
        _LAST_AFFINE = None
        defer = (defer_apply and training and torch.is_grad_enabled() and not relu and res is None and not want_tap
                 and not _BN_FOLD and _DEFER_BN)
        if defer:
            out = _BNNative.apply(xin, gamma, beta, running_mean, running_var, None, decay, eps, False, True, stats,
                                  False, 2 if defer_apply == 2 else 1)
            out._tdl_affine = _LAST_AFFINE
            _LAST_AFFINE = _LAST_BN_SRC = _LAST_RELU_MASK = None
            return out
        if want_tap and training and torch.is_grad_enabled():
            out, tap = _BNNative.apply(xin, gamma, beta, running_mean, running_var, res, decay, eps, relu, training,
                                       stats, True)
        else:
            out = _BNNative.apply(xin, gamma, beta, running_mean, running_var, res, decay, eps, relu,
                                  training, stats if training else None)
        if _LAST_RELU_MASK is not None:
            # lets the next convolution's data-gradient kernel apply this ReLU's backward (see _ConvNative)
            out._tdl_relu_mask = _LAST_RELU_MASK
            _LAST_RELU_MASK = None
        if _LAST_BN_SRC is not None:
            # lets the consumer convolution's data-gradient kernel reduce this BatchNorm's backward sums (_ConvNative)
            out._tdl_bn_src = _LAST_BN_SRC
            _LAST_BN_SRC = None
        if relu:
            out._tdl_relu_out = True
        if want_tap:
            return out, (tap if tap is not None else x)
        return out
    if training:
        y = R.batch_norm_train(x, gamma, beta, running_mean, running_var, decay, eps)
    else:
        y = R.batch_norm_infer(x, gamma, beta, running_mean, running_var, eps)
    if residual is not None:
        y = y + residual
    y = F.relu(y) if relu else y
    return (y, x) if want_tap else y


class _BNConcat(torch.autograd.Function):
    """Several branches that end in BatchNorm (+ReLU) (+ the reference's `_upsample`), concatenated along channels
    WITHOUT a concatenation: every branch normalises (or upsamples) straight into its channel slice of one buffer, and
    in backward every branch reads its slice of the buffer's gradient in place (core/resnet.py:470,485).

    ``meta``: per branch a dict {kind: "bn" | "bn_up" | "up", C, relu, decay, eps, training}; ``tensors``: per branch
    the flattened (x, stats, gamma, beta, moving_mean, moving_variance) for the BatchNorm kinds, (x,) for "up"."""

    @staticmethod
    def forward(ctx, meta, out_hw, *tensors):
        from . import native_ops as O

        ctx.set_materialize_grads(False)
        it = iter(tensors)
        branches = []
        for m in meta:
            if m["kind"] == "up":
                branches.append((m, next(it)))
            else:
                branches.append((m, next(it), next(it), next(it), next(it), next(it), next(it)))
        n = branches[0][1].shape[0]
        oh, ow = out_hw
        ctot = sum(m["C"] for m in meta)
        buf = torch.empty((n, oh, ow, ctot), dtype=torch.bfloat16, device=branches[0][1].device)
        buf2 = buf.view(-1, ctot)
        saved, lo = [], 0
        ctx.params = []
        for b in branches:
            m = b[0]
            C = m["C"]
            sl4, sl2 = buf[..., lo:lo + C], buf2[:, lo:lo + C]
            if m["kind"] == "up":
                O.upsample_fwd(b[1].contiguous(), out_hw, out=sl4)
                _count(1)
                saved.append(None)
                ctx.params.append(None)
            else:
                _, x, stats, gamma, beta, rm, rv = b
                x2 = x.reshape(-1, C)
                if m["training"] and stats is not None:
                    mean, invstd, scale, shift = O.bn_finalize_partials(stats, stats.shape[0], x2.shape[0], gamma, beta,
                                                                        rm, rv, m["decay"], m["eps"])
                    _count(1)
                elif m["training"]:
                    mean, invstd, scale, shift = O.bn_fwd_stats(x2, gamma, beta, rm, rv, m["decay"], m["eps"])
                    _count(2)
                else:
                    scale, shift = O.bn_infer_scale_shift(gamma, beta, rm, rv, m["eps"])
                    mean = invstd = None
                    _count(1)
                if m["kind"] == "bn":
                    O.bn_apply(x2, scale, shift, None, m["relu"], out=sl2)
                    _count(1)
                else:  # normalise the small map, then upsample it into the slice
                    z = O.bn_apply(x2, scale, shift, None, m["relu"]).view(x.shape)
                    O.upsample_fwd(z, out_hw, out=sl4)
                    _count(2)
                saved.append((x2, scale, shift, mean, invstd))
                ctx.params.append((gamma, beta))
            lo += C
        ctx.meta, ctx.saved, ctx.out_hw = meta, saved, out_hw
        ctx.in_shapes = [b[1].shape for b in branches]
        return buf

    @staticmethod
    def backward(ctx, dbuf):
        from . import native_ops as O

        dbuf = dbuf.contiguous()
        ctot = dbuf.shape[-1]
        d2 = dbuf.view(-1, ctot)
        grads, lo = [], 0
        for m, sv, pr, shape in zip(ctx.meta, ctx.saved, ctx.params, ctx.in_shapes):
            C = m["C"]
            if m["kind"] == "up":
                grads.append(O.upsample_bwd(dbuf[..., lo:lo + C], tuple(shape), ctx.out_hw))
                _count(1)
            else:
                if not m["training"]:
                    raise RuntimeError("backward through inference-mode BatchNorm is not supported")
                x2, scale, shift, mean, invstd = sv
                gamma, beta = pr
                direct = _direct_ok(gamma) and _direct_ok(beta)
                if m["kind"] == "bn_up":
                    dz2 = O.upsample_bwd(dbuf[..., lo:lo + C], tuple(shape), ctx.out_hw).reshape(-1, C)
                    _count(1)
                else:
                    dz2 = d2[:, lo:lo + C]  # read in place, row pitch = the whole buffer's channel count
                dx, _, dgamma, dbeta = O.bn_bwd(dz2, x2, None, scale, shift, mean, invstd, m["relu"], False,
                                                acc_dgamma=gamma.grad if direct else None,
                                                acc_dbeta=beta.grad if direct else None)
                _count(4)
                if direct:
                    _notify(gamma)
                    _notify(beta)
                    dgamma = dbeta = None
                grads += [dx.view(shape), None, dgamma if gamma is not None else None,
                          dbeta if beta is not None else None, None, None]
            lo += C
        return (None, None, *grads)


def batch_norm_concat(branches, out_hw):
    """``branches``: list of dicts, each either
         {"x": tensor [N,h,w,C], "bn": BatchNorm module, "relu": bool, "stats": partials or None, "upsample": bool}
       or {"x": tensor [N,h,w,C], "bn": None, "upsample": True}   (an already activated tensor, only resized)
    -> [N, out_h, out_w, sum C]: the channel concatenation of relu?(BN(x)) (resized with the reference's `_upsample` when
    ``upsample``), written slice by slice with no concatenation pass.  CUDA bf16 only; callers keep the plain
    ``torch.cat`` formulation for the torch path."""
    meta, tensors = [], []
    for b in branches:
        x, bn = b["x"], b.get("bn")
        C = x.shape[-1]
        if bn is None:
            meta.append(dict(kind="up", C=C))
            tensors.append(x)
            continue
        meta.append(dict(kind="bn_up" if b.get("upsample") else "bn", C=C, relu=bool(b.get("relu", True)),
                         decay=bn.decay, eps=bn.eps, training=bn.training))
        tensors += [x.contiguous(), b.get("stats") if bn.training else None, bn.gamma, bn.beta, bn.moving_mean,
                    bn.moving_variance]
    return _BNConcat.apply(meta, tuple(out_hw), *tensors)


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
