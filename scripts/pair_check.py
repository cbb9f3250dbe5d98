"""CTA-pair (cta_group::2) convolution kernel: numerics against the single-CTA kernel and fp32, then timings.

    python scripts/pair_check.py [numerics] [timing]
"""
import ctypes
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from tensorflowdistributedlearning_b200 import _native as N
from tensorflowdistributedlearning_b200.ops import native_ops as O

lib = N.lib()
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False


def rel_err(a, b):
    a, b = a.float(), b.float()
    return float((a - b).abs().max() / (b.abs().max() + 1e-6))


def pair(mode):
    N.check(lib.tdl_conv2d_set_pair_mode(ctypes.c_int(mode)), "tdl_conv2d_set_pair_mode")


def ref_conv(x, w, R, S, stride, dil, pads, bias=None, res=None, relu=False):
    K, C = w.shape[0], x.shape[-1]
    w4 = w.float().view(K, R, S, C).permute(0, 3, 1, 2)
    xin = F.pad(x.float().permute(0, 3, 1, 2), (pads[2], pads[3], pads[0], pads[1]))
    y = F.conv2d(xin, w4, None, stride=stride, dilation=dil).permute(0, 2, 3, 1)
    if bias is not None:
        y = y + bias
    if res is not None:
        y = y + res.float()
    return y.relu() if relu else y


def numerics():
    cases = [  # N, H, C, K, R, stride, dil, bias, res, relu
        (4, 14, 128, 512, 1, 1, 1, True, True, True), (3, 13, 72, 136, 1, 1, 1, False, False, False),
        (8, 7, 512, 512, 3, 1, 1, False, False, False), (2, 13, 128, 128, 3, 1, 8, False, False, False),
        (2, 56, 128, 128, 3, 2, 1, False, False, False), (2, 56, 256, 512, 1, 2, 1, False, False, False),
        (3, 14, 256, 256, 3, 1, 1, True, False, True), (4, 28, 128, 512, 1, 1, 1, False, True, False),
        (3, 14, 256, 264, 3, 1, 1, False, True, False), (1, 9, 64, 256, 1, 1, 1, False, False, False),
        (5, 5, 256, 1032, 1, 1, 1, True, True, True), (32, 28, 512, 128, 1, 1, 1, False, False, False)]
    bad = 0
    for (n, h, c, k, r, stride, dil, use_b, use_r, relu) in cases:
        torch.manual_seed(1)
        x = torch.randn(n, h, h, c, device="cuda").bfloat16()
        w = (torch.randn(k, r * r, c, device="cuda") / math.sqrt(r * r * c)).bfloat16()
        b = torch.randn(k, device="cuda") if use_b else None
        P, pt, pb = O.same_pad(h, r, stride, dil)
        res = torch.randn(n, P, P, k, device="cuda").bfloat16() if use_r else None
        outs = {}
        for mode in (0, 1):
            pair(mode)
            outs[mode] = O.conv2d_fprop(x, w, R=r, S=r, stride=stride, dil=dil, pad_h=pt, pad_w=pt, out_hw=(P, P),
                                        bias=b, residual=res, relu=relu)
            torch.cuda.synchronize()
        ref = ref_conv(x, w, r, r, stride, dil, (pt, pb, pt, pb), b, res, relu)
        e1, e0 = rel_err(outs[1], ref), rel_err(outs[0], ref)
        same = bool(torch.equal(outs[0], outs[1]))
        ok = e1 < 2e-2
        bad += not ok
        print(f"fprop {n}x{h}x{h}x{c}->{k} r{r} s{stride} d{dil} b{int(use_b)} res{int(use_r)} relu{int(relu)}: "
              f"pair err {e1:.2e} single err {e0:.2e} bit-equal {same} {'ok' if ok else 'FAIL'}", flush=True)
    # statistics epilogue
    for (n, h, c, k, r, stride) in ((4, 14, 64, 256, 1, 1), (2, 28, 128, 128, 3, 1), (3, 13, 72, 136, 3, 1),
                                    (5, 7, 512, 2048, 1, 1)):
        torch.manual_seed(2)
        x = torch.randn(n, h, h, c, device="cuda").bfloat16()
        w = (torch.randn(k, r * r, c, device="cuda") / math.sqrt(r * r * c)).bfloat16()
        P, pt, pb = O.same_pad(h, r, stride)
        pair(1)
        y, partials, nparts = O.conv2d_fprop_stats(x, w, R=r, S=r, stride=stride, pad_h=pt, pad_w=pt, out_hw=(P, P))
        torch.cuda.synchronize()
        ref = ref_conv(x, w, r, r, stride, 1, (pt, pb, pt, pb))
        yf = y.float().reshape(-1, k)
        tot = partials[:nparts].double().sum(0)
        e = (rel_err(y, ref), rel_err(tot[0], yf.double().sum(0)), rel_err(tot[1], (yf.double() ** 2).sum(0)))
        ok = e[0] < 2e-2 and e[1] < 1e-4 and e[2] < 1e-4
        bad += not ok
        print(f"stats {n}x{h}x{h}x{c}->{k} r{r}: y err {e[0]:.2e} sum err {e[1]:.2e} sumsq err {e[2]:.2e} "
              f"rows {nparts} {'ok' if ok else 'FAIL'}", flush=True)
    # residual on the tensor core + output mask (data gradient of a unit's first conv)
    for (n, h, c, k) in ((4, 14, 64, 256), (3, 13, 128, 1024), (2, 7, 512, 2048)):
        torch.manual_seed(3)
        dy = torch.randn(n, h, h, c, device="cuda").bfloat16()
        wd = (torch.randn(k, 1, c, device="cuda") / math.sqrt(c)).bfloat16()
        res = torch.randn(n, h, h, k, device="cuda").bfloat16()
        mask = torch.randint(0, 256, (n * h * h, k // 8), dtype=torch.uint8, device="cuda")
        outs = {}
        for mode in (0, 1):
            pair(mode)
            outs[mode] = O.conv2d_fprop(dy, wd, R=1, S=1, out_hw=(h, h), residual=res, out_mask=mask)
            torch.cuda.synchronize()
        bits = ((mask.unsqueeze(-1) >> torch.arange(8, device="cuda", dtype=torch.uint8)) & 1).reshape(n, h, h, k)
        ref = ref_conv(dy, wd, 1, 1, 1, 1, (0, 0, 0, 0), None, res) * bits.float()
        e1 = rel_err(outs[1], ref)
        ok = e1 < 2e-2
        bad += not ok
        print(f"resmma+mask {n}x{h}x{h}x{c}->{k}: pair err {e1:.2e} bit-equal {bool(torch.equal(outs[0], outs[1]))} "
              f"{'ok' if ok else 'FAIL'}", flush=True)
    pair(0)
    print("NUMERICS", "OK" if bad == 0 else f"FAILED ({bad})", flush=True)
    return bad


def timed(fn, iters=20, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    ts = []
    for _ in range(iters):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def timing(out="gpurun_out/pair_timing.txt"):
    shapes = [  # ResNet-50 layers at batch 256: N, H, C, K, R, stride
        (256, 56, 64, 256, 1, 1), (256, 56, 256, 64, 1, 1), (256, 28, 128, 128, 3, 1), (256, 28, 512, 128, 1, 1),
        (256, 28, 128, 512, 1, 1), (256, 14, 256, 256, 3, 1), (256, 14, 1024, 256, 1, 1), (256, 14, 256, 1024, 1, 1),
        (256, 7, 512, 512, 3, 1), (256, 7, 2048, 512, 1, 1), (256, 7, 512, 2048, 1, 1), (256, 56, 128, 128, 3, 2),
        (256, 56, 256, 512, 1, 2), (256, 28, 256, 256, 3, 2), (256, 14, 512, 512, 3, 2)]
    lines = ["# fprop, us (median of 20, L2 flushed): single-CTA MMAs vs CTA pairs (cta_group::2); TFLOP/s of the pair kernel",
             "# shape [N, H=W, Cin, Cout, R=S, stride]      single     pair   speedup   TFLOP/s"]
    for (n, h, c, k, r, stride) in shapes:
        x = torch.randn(n, h, h, c, device="cuda").bfloat16()
        w = (torch.randn(k, r * r, c, device="cuda") / math.sqrt(r * r * c)).bfloat16()
        P, pt, pb = O.same_pad(h, r, stride)
        t = {}
        for mode in (0, 1):
            pair(mode)
            t[mode] = timed(lambda: O.conv2d_fprop(x, w, R=r, S=r, stride=stride, pad_h=pt, pad_w=pt, out_hw=(P, P)))
        flops = 2.0 * n * P * P * k * c * r * r
        lines.append(f"  {str([n, h, c, k, r, stride]):34s} {t[0]:8.1f} {t[1]:8.1f} {t[0] / t[1]:8.2f} {flops / t[1] / 1e6:9.0f}")
        print(lines[-1], flush=True)
    pair(0)
    os.makedirs(os.path.dirname(out), exist_ok=True)
    open(out, "w").write("\n".join(lines) + "\n")


if __name__ == "__main__":
    which = sys.argv[1:] or ["numerics", "timing"]
    rc = 0
    if "numerics" in which:
        rc = numerics()
    if "timing" in which and rc == 0:
        timing()
    sys.exit(1 if rc else 0)
