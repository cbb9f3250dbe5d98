"""Numerics + timing probes for the native kernels, one group per process.

Usage (on a B200):  python scripts/gpu_check.py <group> [...]   or   python scripts/gpu_check.py all
Each group runs in its own subprocess under a timeout so that a trapped kernel (which poisons the
CUDA context) cannot take the other groups down with it.  Results are appended as JSON lines to
gpurun_out/gpu_check.jsonl.
"""
from __future__ import annotations

import json
import os
import subprocess
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
os.makedirs(OUT, exist_ok=True)
LOG = os.path.join(OUT, "gpu_check.jsonl")


def emit(**kw):
    kw["t"] = time.time()
    line = json.dumps(kw)
    print(line, flush=True)
    with open(LOG, "a") as f:
        f.write(line + "\n")


def rel_err(a, b):
    import torch

    a = a.float()
    b = b.float()
    return float((a - b).abs().max() / (b.abs().max() + 1e-6))


def time_cuda(fn, iters=20, warmup=3):
    import torch

    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    ts = []
    for _ in range(iters):
        flush.zero_()
        s = torch.cuda.Event(enable_timing=True)
        e = torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


def ref_conv(x, w_ohwi, R, S, stride, dil, pads, bias=None, residual=None, relu=False):
    """fp32 reference on the same bf16-rounded inputs.  x NHWC, w [K, R*S, C]; pads=(t,b,l,r)."""
    import torch
    import torch.nn.functional as F

    K = w_ohwi.shape[0]
    C = x.shape[-1]
    w = w_ohwi.float().view(K, R, S, C).permute(0, 3, 1, 2).contiguous()
    xin = x.float().permute(0, 3, 1, 2)
    xin = F.pad(xin, (pads[2], pads[3], pads[0], pads[1]))
    y = F.conv2d(xin, w, None, stride=stride, dilation=dil)
    y = y.permute(0, 2, 3, 1)
    if bias is not None:
        y = y + bias
    if residual is not None:
        y = y + residual.float()
    if relu:
        y = y.relu()
    return y.contiguous()


def conv_case(name, N, H, W, C, K, R, S, stride=1, dil=1, same=True, bias=False, res=False, relu=False):
    import torch
    from tensorflowdistributedlearning_b200.ops import native_ops as O

    torch.manual_seed(0)
    x = torch.randn(N, H, W, C, device="cuda").bfloat16()
    w = (torch.randn(K, R * S, C, device="cuda") / (R * S * C) ** 0.5).bfloat16()
    if same:
        P, pt, pb = O.same_pad(H, R, stride, dil)
        Q, pl, pr = O.same_pad(W, S, stride, dil)
    else:
        pt = pb = pl = pr = 0
        P = (H - ((R - 1) * dil + 1)) // stride + 1
        Q = (W - ((S - 1) * dil + 1)) // stride + 1
    b = torch.randn(K, device="cuda") if bias else None
    r = torch.randn(N, P, Q, K, device="cuda").bfloat16() if res else None
    y = O.conv2d_fprop(x, w, R=R, S=S, stride=stride, dil=dil, pad_h=pt, pad_w=pl, out_hw=(P, Q), bias=b,
                       residual=r, relu=relu)
    torch.cuda.synchronize()
    yr = ref_conv(x, w, R, S, stride, dil, (pt, pb, pl, pr), b, r, relu)
    err = rel_err(y, yr)
    emit(group="conv_fprop", case=name, shape=[N, H, W, C, K, R, S, stride, dil], rel_err=err, ok=err < 2e-2)
    return x, w, (P, Q, pt, pl)


def group_conv_fprop_basic():
    conv_case("gemm_64_64", 2, 16, 16, 64, 64, 1, 1)
    conv_case("gemm_256_128", 2, 28, 28, 256, 128, 1, 1)
    conv_case("gemm_128_512_bias_res_relu", 4, 14, 14, 128, 512, 1, 1, bias=True, res=True, relu=True)
    conv_case("gemm_ragged_c72_k40", 3, 13, 13, 72, 40, 1, 1)
    conv_case("c3x3_64", 2, 16, 16, 64, 64, 3, 3)
    conv_case("c3x3_56", 2, 56, 56, 64, 64, 3, 3)
    conv_case("c3x3_14_256", 4, 14, 14, 256, 256, 3, 3)
    conv_case("c3x3_7_512_multiimg", 8, 7, 7, 512, 512, 3, 3)
    conv_case("c3x3_13_dil2", 2, 13, 13, 128, 128, 3, 3, dil=2)
    conv_case("c3x3_13_dil8", 2, 13, 13, 64, 64, 3, 3, dil=8)


def group_conv_fprop_strided():
    conv_case("c1x1_s2", 2, 56, 56, 256, 512, 1, 1, stride=2)
    conv_case("c3x3_s2_even", 2, 56, 56, 128, 128, 3, 3, stride=2)
    conv_case("c3x3_s2_odd", 2, 51, 51, 64, 64, 3, 3, stride=2)
    conv_case("c3x3_s2_26", 2, 26, 26, 128, 128, 3, 3, stride=2)


def group_conv_wgrad():
    import torch
    import torch.nn.functional as F
    from tensorflowdistributedlearning_b200.ops import native_ops as O

    def case(name, N, H, W, C, K, R, S, stride=1, dil=1):
        torch.manual_seed(1)
        x = torch.randn(N, H, W, C, device="cuda").bfloat16()
        P, pt, pb = O.same_pad(H, R, stride, dil)
        Q, pl, pr = O.same_pad(W, S, stride, dil)
        dy = torch.randn(N, P, Q, K, device="cuda").bfloat16()
        dw = torch.zeros(K, R * S, C, device="cuda")
        O.conv2d_wgrad(x, dy, dw, R=R, S=S, stride=stride, dil=dil, pad_h=pt, pad_w=pl)
        torch.cuda.synchronize()
        xin = F.pad(x.float().permute(0, 3, 1, 2), (pl, pr, pt, pb)).requires_grad_(False)
        wr = torch.zeros(K, C, R, S, device="cuda", requires_grad=True)
        yr = F.conv2d(xin, wr, None, stride=stride, dilation=dil)
        (gw,) = torch.autograd.grad(yr, wr, dy.float().permute(0, 3, 1, 2))
        ref = gw.permute(0, 2, 3, 1).reshape(K, R * S, C)
        err = rel_err(dw, ref)
        emit(group="conv_wgrad", case=name, shape=[N, H, W, C, K, R, S, stride, dil], rel_err=err, ok=err < 2e-2)

    case("w_gemm_64_64", 2, 16, 16, 64, 64, 1, 1)
    case("w_gemm_256_128", 2, 28, 28, 256, 128, 1, 1)
    case("w_gemm_ragged", 3, 13, 13, 72, 40, 1, 1)
    case("w_3x3_64", 2, 16, 16, 64, 64, 3, 3)
    case("w_3x3_14_256", 4, 14, 14, 256, 256, 3, 3)
    case("w_3x3_7_multi", 8, 7, 7, 128, 128, 3, 3)
    case("w_3x3_dil2", 2, 13, 13, 128, 128, 3, 3, dil=2)
    case("w_3x3_s2", 2, 56, 56, 64, 128, 3, 3, stride=2)
    case("w_1x1_s2", 2, 28, 28, 128, 256, 1, 1, stride=2)


def group_conv_dgrad():
    import torch
    import torch.nn.functional as F
    from tensorflowdistributedlearning_b200.ops import native_ops as O

    def case(name, N, H, W, C, K, R, S, dil=1):
        torch.manual_seed(2)
        P, pt, pb = O.same_pad(H, R, 1, dil)
        Q, pl, pr = O.same_pad(W, S, 1, dil)
        wm = torch.randn(K, R * S, C, device="cuda") / (R * S * C) ** 0.5
        wf = torch.empty(K, R * S, C, device="cuda", dtype=torch.bfloat16)
        wd = torch.empty(C, R * S, K, device="cuda", dtype=torch.bfloat16)
        O.conv2d_weight_prep(wm, wf, wd, K, R * S, C)
        dy = torch.randn(N, P, Q, K, device="cuda").bfloat16()
        dx = O.conv2d_fprop(dy, wd, R=R, S=S, stride=1, dil=dil, pad_h=(R - 1) * dil - pt, pad_w=(S - 1) * dil - pl,
                            out_hw=(H, W))
        torch.cuda.synchronize()
        xr = torch.zeros(N, C, H + pt + pb, W + pl + pr, device="cuda", requires_grad=True)
        w4 = wf.float().view(K, R, S, C).permute(0, 3, 1, 2).contiguous()
        yr = F.conv2d(xr, w4, None, stride=1, dilation=dil)
        (gx,) = torch.autograd.grad(yr, xr, dy.float().permute(0, 3, 1, 2))
        ref = gx[:, :, pt:pt + H, pl:pl + W].permute(0, 2, 3, 1)
        err = rel_err(dx, ref)
        emit(group="conv_dgrad", case=name, rel_err=err, ok=err < 2e-2,
             wf_ok=bool(torch.equal(wf, wm.bfloat16())))

    case("d_1x1", 2, 14, 14, 256, 64, 1, 1)
    case("d_3x3", 2, 14, 14, 128, 128, 3, 3)
    case("d_3x3_dil4", 2, 13, 13, 64, 128, 3, 3, dil=4)


def group_bn():
    import torch
    from tensorflowdistributedlearning_b200.ops import native_ops as O

    for (rows, C, relu, res) in [(4096, 64, True, False), (5000, 256, True, True), (777, 2048, False, False),
                                 (100352, 64, True, False)]:
        torch.manual_seed(3)
        x = (torch.randn(rows, C, device="cuda") * 2 + 0.5).bfloat16()
        gamma = torch.rand(C, device="cuda") + 0.5
        beta = torch.randn(C, device="cuda") * 0.1
        rm = torch.zeros(C, device="cuda")
        rv = torch.ones(C, device="cuda")
        r = torch.randn(rows, C, device="cuda").bfloat16() if res else None
        mean, invstd, scale, shift = O.bn_fwd_stats(x, gamma, beta, rm, rv, 0.99, 1e-3)
        z = O.bn_apply(x, scale, shift, r, relu)
        xf = x.float()
        m_ref = xf.mean(0)
        v_ref = xf.var(0, unbiased=False)
        zr = (xf - m_ref) / torch.sqrt(v_ref + 1e-3) * gamma + beta
        if res:
            zr = zr + r.float()
        if relu:
            zr = zr.relu()
        e_mean = rel_err(mean, m_ref)
        e_var = rel_err(1 / invstd ** 2 - 1e-3, v_ref)
        e_z = rel_err(z, zr)
        e_rm = rel_err(rm, 0.01 * m_ref)
        e_rv = rel_err(rv, 0.99 + 0.01 * xf.var(0, unbiased=True))
        # backward
        dz = torch.randn(rows, C, device="cuda").bfloat16()
        dx, dres, dg, db = O.bn_bwd(dz, x, z if res else None, scale, shift, mean, invstd, relu, res)
        xr = xf.clone().requires_grad_(True)
        g_ = gamma.clone().requires_grad_(True)
        b_ = beta.clone().requires_grad_(True)
        rr = r.float().clone().requires_grad_(True) if res else None
        mm = xr.mean(0)
        vv = xr.var(0, unbiased=False)
        out = (xr - mm) / torch.sqrt(vv + 1e-3) * g_ + b_
        if res:
            out = out + rr
        if relu:
            # use the kernel's own mask (bf16 rounding makes near-zero entries ambiguous)
            out = out * (z.float() > 0).float()
        out.backward(dz.float())
        e_dx = rel_err(dx, xr.grad)
        e_dg = rel_err(dg, g_.grad)
        e_db = rel_err(db, b_.grad)
        e_dr = rel_err(dres, rr.grad) if res else 0.0
        ok = max(e_mean, e_var, e_z, e_rm, e_rv, e_dx, e_dg, e_db, e_dr) < 2e-2
        emit(group="bn", rows=rows, C=C, relu=relu, res=res, e_mean=e_mean, e_var=e_var, e_z=e_z, e_rm=e_rm,
             e_rv=e_rv, e_dx=e_dx, e_dg=e_dg, e_db=e_db, e_dr=e_dr, ok=ok)


def group_optim():
    import torch
    from tensorflowdistributedlearning_b200.ops import native_ops as O

    n = 1_000_003
    torch.manual_seed(4)
    p = torch.randn(n, device="cuda")
    g = torch.randn(n, device="cuda")
    for gd in (torch.float32, torch.bfloat16):
        p1 = p.clone(); m = torch.zeros(n, device="cuda"); sh = torch.empty(n, device="cuda", dtype=torch.bfloat16)
        step = torch.zeros(1, dtype=torch.int64, device="cuda"); lr_state = torch.zeros(2, device="cuda")
        pr = p.clone(); mr = torch.zeros(n, device="cuda")
        for it in range(3):
            gg = g.to(gd).clone()
            O.lr_schedule_step(step, lr_state, 0.1, 0.5, 10.0)
            O.sgd_momentum_flat(p1, gg, m, sh, lr_state, 0.9, 1e-4, 0.5, True)
            lr = 0.1 * 0.5 ** (it / 10.0)
            gr = g.to(gd).float() * 0.5 + 1e-4 * pr
            mr = 0.9 * mr + gr
            pr = pr - lr * mr
            assert float(gg.float().abs().max()) == 0.0
        emit(group="optim", kind="sgdm", grad=str(gd), e_p=rel_err(p1, pr), e_sh=rel_err(sh, pr), step=int(step.item()),
             ok=rel_err(p1, pr) < 1e-5 and int(step.item()) == 3)
        p1 = p.clone(); m1 = torch.zeros(n, device="cuda"); m2 = torch.zeros(n, device="cuda")
        step.zero_()
        pr = p.clone(); a = torch.zeros(n, device="cuda"); b = torch.zeros(n, device="cuda")
        for it in range(3):
            gg = g.to(gd).clone()
            O.lr_schedule_step(step, lr_state, 1e-3, 0.5, 10000.0)
            O.adam_flat(p1, gg, m1, m2, None, lr_state, 0.9, 0.999, 1e-8, 1.0, True)
            t = it + 1
            lr = 1e-3 * 0.5 ** (it / 10000.0)
            lr_t = lr * (1 - 0.999 ** t) ** 0.5 / (1 - 0.9 ** t)
            gr = g.to(gd).float()
            a = 0.9 * a + 0.1 * gr
            b = 0.999 * b + 0.001 * gr * gr
            pr = pr - lr_t * a / (b.sqrt() + 1e-8)
        emit(group="optim", kind="adam", grad=str(gd), e_p=rel_err(p1, pr), ok=rel_err(p1, pr) < 1e-4)


def group_loss():
    import torch
    from tensorflowdistributedlearning_b200.ops import native_ops as O
    from tensorflowdistributedlearning_b200.ops import reference as R

    torch.manual_seed(5)
    for (B, P) in [(4, 10201), (3, 1000), (2, 16384), (5, 77)]:
        logits = torch.randn(B, P, device="cuda") * 2
        labels = (torch.rand(B, P, device="cuda") > 0.7).float()
        if B > 1:
            labels[1] = 0  # empty mask image
        loss, dl = O.lovasz_hinge(logits, labels, True, 1.0 / B)
        lg = logits.double().cpu().requires_grad_(True)
        ref = R.lovasz_hinge_per_image(lg, labels.double().cpu())
        ref.mean().backward()
        e_l = rel_err(loss.cpu(), ref.detach())
        e_g = rel_err(dl.cpu(), lg.grad)
        iou, acc = O.seg_metrics(logits, labels, 0.0, True)
        riou, racc = R.seg_metrics(logits.cpu(), labels.cpu(), 0.5, True)
        e_i = rel_err(iou.cpu(), riou)
        e_a = rel_err(acc.cpu(), racc)
        emit(group="loss", kind="lovasz", B=B, P=P, e_loss=e_l, e_grad=e_g, e_iou=e_i, e_acc=e_a,
             ok=max(e_l, e_g, e_i, e_a) < 1e-4)
    for (B, C) in [(256, 1000), (7, 10), (33, 21841)]:
        logits = torch.randn(B, C, device="cuda") * 3
        labels = torch.randint(0, C, (B,), device="cuda")
        loss, dl, t1, t5 = O.softmax_ce(logits, labels, True, 1.0 / B, 5)
        lr = logits.clone().requires_grad_(True)
        ref = torch.nn.functional.cross_entropy(lr, labels, reduction="none")
        ref.mean().backward()
        top5 = (logits.topk(min(5, C), dim=1).indices == labels[:, None]).any(1).float()
        top1 = (logits.argmax(1) == labels).float()
        emit(group="loss", kind="softmax_ce", B=B, C=C, e_loss=rel_err(loss, ref.detach()), e_grad=rel_err(dl, lr.grad),
             top1=bool(torch.equal(t1, top1)), top5=bool(torch.equal(t5, top5)),
             ok=rel_err(loss, ref.detach()) < 1e-4 and rel_err(dl, lr.grad) < 1e-3)


def group_perf():
    """ResNet-50 layer shapes at batch 256: ours vs cuDNN (torch channels_last bf16)."""
    import torch
    import torch.nn.functional as F
    from tensorflowdistributedlearning_b200.ops import native_ops as O

    shapes = [  # N, H, C, K, R, stride
        (256, 56, 64, 64, 1, 1), (256, 56, 64, 64, 3, 1), (256, 56, 64, 256, 1, 1), (256, 56, 256, 64, 1, 1),
        (256, 28, 128, 128, 3, 1), (256, 28, 512, 128, 1, 1), (256, 28, 128, 512, 1, 1),
        (256, 14, 256, 256, 3, 1), (256, 14, 1024, 256, 1, 1), (256, 14, 256, 1024, 1, 1),
        (256, 7, 512, 512, 3, 1), (256, 7, 2048, 512, 1, 1), (256, 7, 512, 2048, 1, 1),
        (256, 56, 128, 128, 3, 2), (256, 56, 256, 512, 1, 2),
    ]
    for (N, H, C, K, R, stride) in shapes:
        try:
            x = torch.randn(N, H, H, C, device="cuda").bfloat16()
            w = (torch.randn(K, R * R, C, device="cuda") * 0.05).bfloat16()
            P, pt, pb = O.same_pad(H, R, stride)
            pad_sym = (R - 1) // 2
            t_fwd = time_cuda(lambda: O.conv2d_fprop(x, w, R=R, S=R, stride=stride, pad_h=pad_sym, pad_w=pad_sym,
                                                     out_hw=(P, P)))
            xc = x.permute(0, 3, 1, 2)  # channels_last view
            wc = w.view(K, R, R, C).permute(0, 3, 1, 2)
            t_cudnn = time_cuda(lambda: F.conv2d(xc, wc, None, stride=stride, padding=pad_sym))
            dy = torch.randn(N, P, P, K, device="cuda").bfloat16()
            dw = torch.zeros(K, R * R, C, device="cuda")
            t_wg = time_cuda(lambda: O.conv2d_wgrad(x, dy, dw, R=R, S=R, stride=stride, pad_h=pad_sym, pad_w=pad_sym))
            dyc = dy.permute(0, 3, 1, 2)
            t_wg_cudnn = time_cuda(lambda: torch.ops.aten.convolution_backward(
                dyc, xc, wc, None, [stride, stride], [pad_sym, pad_sym], [1, 1], False, [0, 0], 1,
                [False, True, False]))
            flops = 2.0 * N * P * P * K * C * R * R
            emit(group="perf", shape=[N, H, C, K, R, stride], fwd_ms=t_fwd, cudnn_fwd_ms=t_cudnn,
                 fwd_tflops=flops / t_fwd / 1e9, cudnn_tflops=flops / t_cudnn / 1e9, wgrad_ms=t_wg,
                 cudnn_wgrad_ms=t_wg_cudnn, wgrad_tflops=flops / t_wg / 1e9)
        except Exception as e:  # noqa: BLE001
            emit(group="perf", shape=[N, H, C, K, R, stride], error=repr(e)[:300])
    # BN apply / stats bandwidth
    rows, C = 256 * 56 * 56, 256
    x = torch.randn(rows, C, device="cuda").bfloat16()
    gamma = torch.ones(C, device="cuda"); beta = torch.zeros(C, device="cuda")
    rm = torch.zeros(C, device="cuda"); rv = torch.ones(C, device="cuda")
    t_stats = time_cuda(lambda: O.bn_fwd_stats(x, gamma, beta, rm, rv, 0.99, 1e-3))
    mean, invstd, scale, shift = O.bn_fwd_stats(x, gamma, beta, rm, rv, 0.99, 1e-3)
    t_apply = time_cuda(lambda: O.bn_apply(x, scale, shift, None, True))
    dz = torch.randn(rows, C, device="cuda").bfloat16()
    t_bwd = time_cuda(lambda: O.bn_bwd(dz, x, None, scale, shift, mean, invstd, True, False))
    gb = rows * C * 2 / 1e9
    emit(group="perf", kind="bn", rows=rows, C=C, stats_ms=t_stats, stats_gbs=gb / t_stats * 1e3, apply_ms=t_apply,
         apply_gbs=2 * gb / t_apply * 1e3, bwd_ms=t_bwd, bwd_gbs=5 * gb / t_bwd * 1e3)


def group_spatial():
    import torch
    import torch.nn.functional as F
    from tensorflowdistributedlearning_b200.ops import functional as TFn
    from tensorflowdistributedlearning_b200.ops import pooling, depthwise, resize
    from tensorflowdistributedlearning_b200.ops import reference as R

    def run_pair(name, fn, inputs, tol=2e-2):
        """fn(*tensors) -> output; compares native (bf16 cuda) against the torch oracle in fp32, fwd and bwd."""
        torch.manual_seed(7)
        nat_in = [t.clone().requires_grad_(t.is_floating_point()) for t in inputs]
        ref_in = [t.float().clone().requires_grad_(t.is_floating_point()) for t in inputs]
        TFn.force_reference(False)
        y = fn(*nat_in)
        g = torch.randn_like(y)
        y.backward(g)
        TFn.force_reference(True)
        yr = fn(*ref_in)
        yr.backward(g.float())
        TFn.force_reference(False)
        errs = {"fwd": rel_err(y, yr)}
        for i, (a, b) in enumerate(zip(nat_in, ref_in)):
            if a.grad is not None:
                errs[f"g{i}"] = rel_err(a.grad, b.grad)
        emit(group="spatial", case=name, ok=max(errs.values()) < tol, **errs)

    x = torch.randn(4, 51, 51, 128, device="cuda").bfloat16()
    run_pair("maxpool_same_51", lambda t: pooling.max_pool_same(t, 3, 2), [x])
    x = torch.randn(4, 112, 112, 64, device="cuda").bfloat16().relu()
    run_pair("maxpool_sym_112_ties", lambda t: pooling.max_pool(t, 3, 2, (1, 1, 1, 1)), [x])
    x = torch.randn(8, 7, 7, 2048, device="cuda").bfloat16()
    run_pair("gap", lambda t: pooling.global_avg_pool(t, True), [x])
    for (H, C, st, dil) in [(13, 1024, 1, 2), (13, 1024, 1, 8), (38, 728, 1, 1), (75, 256, 2, 1), (19, 728, 2, 1)]:
        x = torch.randn(2, H, H, C, device="cuda").bfloat16()
        w = (torch.randn(3, 3, C, device="cuda") * 0.3)
        b = torch.randn(C, device="cuda") * 0.1
        run_pair(f"depthwise_{H}_{C}_s{st}_d{dil}",
                 lambda t, ww, bb: depthwise.depthwise_conv2d(t, ww, bb, st, dil, "SAME", True), [x, w, b], tol=3e-2)
    for (h, C, out, dt) in [(1, 256, 13, torch.bfloat16), (13, 256, 26, torch.bfloat16), (26, 1, 101, torch.float32)]:
        x = torch.randn(2, h, h, C, device="cuda").to(dt)
        run_pair(f"upsample_{h}_{out}", lambda t: resize.upsample(t, (out, out)), [x], tol=2e-2)


def group_p2p_emulate():
    """All 'ranks' on one GPU, launched one after another with barriers disabled: checks the reduce/optimizer math."""
    import torch
    from tensorflowdistributedlearning_b200.parallel.p2p import SymmetricBuffer, P2PComm

    for world in (2, 4, 8):
        for kind, n in ((1, 1_000_003 // 4 * 4 + 3), (2, 65_536 + 5), (2, 3_000_000)):
            torch.manual_seed(world * 10 + kind)
            gbuf = SymmetricBuffer(((n + 7) // 8 * 8) * 4, emulate_world=world)
            pbuf = SymmetricBuffer(((n + 7) // 8 * 8) * 4, emulate_world=world)
            grads = [torch.randn(n, device="cuda") for _ in range(world)]
            p0 = torch.randn(n, device="cuda")
            for r in range(world):
                gbuf.tensor(torch.float32, r)[:n].copy_(grads[r])
                pbuf.tensor(torch.float32, r)[:n].copy_(p0)
            lr_state = torch.tensor([0.1, 0.05], device="cuda")
            res = {}
            for mode in (1, 2):
                for r in range(world):
                    pbuf.tensor(torch.float32, r)[:n].copy_(p0)
                m1s = [torch.zeros(n + 8, device="cuda") for _ in range(world)]
                m2s = [torch.zeros(n + 8, device="cuda") for _ in range(world)]
                comms = [P2PComm(gbuf, pbuf, emulate_rank=r) for r in range(world)]
                for r in range(world):
                    comms[r].fused_step(0, n, kind=kind, grad_is_bf16=False, zero_grad=False, m1=m1s[r], m2=m2s[r],
                                        lr_state=lr_state, a=0.9, b=(1e-4 if kind == 1 else 0.999), c=1e-8, mode=mode)
                torch.cuda.synchronize()
                g = sum(grads) / world
                if kind == 1:
                    gg = g + 1e-4 * p0
                    ref = p0 - 0.1 * gg
                else:
                    m = 0.1 * g
                    v = 0.001 * g * g
                    ref = p0 - 0.05 * m / (v.sqrt() + 1e-8)
                errs = [rel_err(pbuf.tensor(torch.float32, r)[:n], ref) for r in range(world)]
                same = all(torch.equal(pbuf.tensor(torch.float32, r)[:n], pbuf.tensor(torch.float32, 0)[:n])
                           for r in range(world))
                res[f"mode{mode}"] = max(errs)
                res[f"mode{mode}_mirrors_identical"] = same
                for c in comms:
                    c.close()
            emit(group="p2p_emulate", world=world, kind=kind, n=n, ok=max(res["mode1"], res["mode2"]) < 1e-5 and
                 res["mode1_mirrors_identical"] and res["mode2_mirrors_identical"], **res)
            gbuf.close(); pbuf.close()


def group_model():
    """ResNet-50 (small input) one training step: native kernels vs the torch-op path on identical weights."""
    import torch
    from tensorflowdistributedlearning_b200 import models
    from tensorflowdistributedlearning_b200.ops import functional as TFn

    torch.manual_seed(11)
    m = models.resnet50(num_classes=1000).cuda()
    for mod in m.modules():  # make the zero-init gammas non-trivial so every branch carries signal
        if hasattr(mod, "gamma") and mod.gamma is not None:
            torch.nn.init.uniform_(mod.gamma, 0.5, 1.5)
    x = torch.randn(16, 64, 64, 3, device="cuda").bfloat16()
    y = torch.randint(0, 1000, (16,), device="cuda")
    out = {}
    for name, flag in (("native", False), ("torch", True)):
        TFn.force_reference(flag)
        m.zero_grad(set_to_none=True)
        for mod in m.modules():
            if hasattr(mod, "moving_mean"):
                mod.moving_mean.zero_(); mod.moving_variance.fill_(1.0)
        logits = m(x)
        loss, t1, t5 = TFn.softmax_cross_entropy(logits.float(), y)
        loss.backward()
        out[name] = (float(loss), {n: p.grad.float().clone() for n, p in m.named_parameters()},
                     logits.float().detach().clone())
    TFn.force_reference(False)
    gn, gt = out["native"][1], out["torch"][1]
    worst, worst_name = 0.0, ""
    import math
    cos_min = 1.0
    for n in gn:
        a, b = gn[n].flatten(), gt[n].flatten()
        cos = float(torch.dot(a, b) / (a.norm() * b.norm() + 1e-12))
        if cos < cos_min:
            cos_min, worst_name = cos, n
    emit(group="model", loss_native=out["native"][0], loss_torch=out["torch"][0],
         logits_rel_err=rel_err(out["native"][2], out["torch"][2]), min_grad_cosine=cos_min, worst=worst_name,
         ok=abs(out["native"][0] - out["torch"][0]) < 0.05 and cos_min > 0.9)


def group_perf_depthwise():
    """Depthwise 3x3 kernels on Xception-41 shapes (batch 64): achieved GB/s of algorithmic traffic."""
    import torch
    from tensorflowdistributedlearning_b200.ops import native_ops as O

    for (H, C) in [(19, 728), (38, 728), (75, 256), (150, 128)]:
        x = torch.randn(64, H, H, C, device="cuda").bfloat16()
        w = torch.randn(9, C, device="cuda")
        dy = torch.randn(64, H, H, C, device="cuda").bfloat16()
        gb = x.numel() * 2 / 1e9
        t_f = time_cuda(lambda: O.depthwise3x3_fwd(x, w, None, None, 1, 1, 1, 1, (H, H), True, False))
        t_d = time_cuda(lambda: O.depthwise3x3_fwd(dy, w, None, x, 1, 1, 1, 1, (H, H), False, False))
        t_w = time_cuda(lambda: O.depthwise3x3_wgrad(x, dy, None, 1, 1, 1, 1, True, False))
        emit(group="perf_depthwise", H=H, C=C, fwd_us=t_f * 1e3, fwd_gbs=2 * gb / t_f * 1e3, dgrad_us=t_d * 1e3,
             dgrad_gbs=3 * gb / t_d * 1e3, wgrad_us=t_w * 1e3, wgrad_gbs=2 * gb / t_w * 1e3)


def group_perf_stem():
    """First-layer kernels (pack, weight prep, tcgen05 stem with fused statistics) vs cuDNN on the ResNet-50 /
    Xception-41 / reference roots; max-pool 3x3 s2 fwd/bwd on the ResNet-50 shape."""
    import torch
    import torch.nn.functional as F
    from tensorflowdistributedlearning_b200.ops import native_ops as O
    from tensorflowdistributedlearning_b200.ops import functional as Fn

    for (N, H, C, K, R, padding) in [(256, 224, 3, 64, 7, (3, 3, 3, 3)), (64, 299, 3, 32, 3, (1, 1, 1, 1)),
                                     (32, 101, 2, 64, 3, "SAME")]:
        x = torch.randn(N, H, H, C, device="cuda").bfloat16()
        w = torch.randn(K, R, R, C, device="cuda") / (R * R * C) ** 0.5
        P, Q, pt, pb, pl, pr = Fn.conv_pads(H, H, R, R, 2, 1, padding)
        planes = O.conv_stem_pack(x)
        wp = O.conv_stem_weight_prep(w, pt, pl)
        t_pack = time_cuda(lambda: O.conv_stem_pack(x))
        t_prep = time_cuda(lambda: O.conv_stem_weight_prep(w, pt, pl))
        t_conv = time_cuda(lambda: O.conv_stem_fprop(planes, wp, (H, H), K, R, R, pt, pl, (P, Q), True))
        xin = F.pad(x.permute(0, 3, 1, 2), (pl, pr, pt, pb))
        w4 = w.bfloat16().permute(0, 3, 1, 2)
        t_cudnn = time_cuda(lambda: F.conv2d(xin, w4, None, stride=2))
        out_gb = N * P * Q * K * 2 / 1e9
        emit(group="perf_stem", N=N, H=H, C=C, K=K, R=R, pack_us=t_pack * 1e3, wprep_us=t_prep * 1e3,
             stem_us=t_conv * 1e3, stem_out_gbs=out_gb / t_conv * 1e3, cudnn_fprop_us=t_cudnn * 1e3)
    x = torch.randn(256, 112, 112, 64, device="cuda").bfloat16()
    y, idx = O.maxpool_fwd(x, 3, 2, 0, 0, (56, 56))
    dy = torch.randn_like(y)
    t_f = time_cuda(lambda: O.maxpool_fwd(x, 3, 2, 0, 0, (56, 56)))
    t_b = time_cuda(lambda: O.maxpool_bwd(dy, idx, tuple(x.shape), 3, 2, 0, 0))
    gb = x.numel() * 2 / 1e9
    emit(group="perf_stem", op="maxpool3x3s2", fwd_us=t_f * 1e3, fwd_gbs=(gb * 1.25 + gb / 8) / t_f * 1e3,
         bwd_us=t_b * 1e3, bwd_gbs=(gb * 1.25 + gb / 8) / t_b * 1e3)


GROUPS = {
    "conv_fprop_basic": group_conv_fprop_basic,
    "conv_fprop_strided": group_conv_fprop_strided,
    "conv_wgrad": group_conv_wgrad,
    "conv_dgrad": group_conv_dgrad,
    "bn": group_bn,
    "optim": group_optim,
    "loss": group_loss,
    "perf": group_perf,
    "spatial": group_spatial,
    "p2p_emulate": group_p2p_emulate,
    "model": group_model,
    "perf_depthwise": group_perf_depthwise,
    "perf_stem": group_perf_stem,
}


def main():
    args = sys.argv[1:] or ["all"]
    if args == ["all"]:
        rc_all = 0
        for g in GROUPS:
            try:
                r = subprocess.run([sys.executable, os.path.abspath(__file__), g], timeout=420)
                rc = r.returncode
            except subprocess.TimeoutExpired:
                rc = -9
            emit(group=g, event="group_exit", rc=rc)
            rc_all |= int(rc != 0)
        sys.exit(rc_all)
    for g in args:
        try:
            GROUPS[g]()
        except Exception as e:  # noqa: BLE001
            import traceback

            emit(group=g, event="exception", error=repr(e)[:500], tb=traceback.format_exc()[-1500:])
            sys.exit(1)


if __name__ == "__main__":
    main()
