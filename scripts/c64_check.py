"""64-channel 3x3 layers (ResNet-50 stage 1, the reference network's root): the resident-weight halo loop of the forward /
data-gradient kernel and the row mode of the weight gradient, each against the loop it replaces.
CUDA events, L2 flushed between iterations, median of 20.

    python scripts/c64_check.py [--out gpurun_out/c64_timing.txt]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from tensorflowdistributedlearning_b200 import _native
from tensorflowdistributedlearning_b200.ops import native_ops as O

ap = argparse.ArgumentParser()
ap.add_argument("--out", default="gpurun_out/c64_timing.txt")
a = ap.parse_args()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
lib = _native.lib()


def timed(fn, iters=20, warmup=4):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


def rel(a, b):
    return float((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-20))


lines = ["# 3x3 stride-1 layers with <= 64 input channels at batch 256: us per call (median of 20, L2 flushed)",
         "# fprop: tap-shifted loop vs halo loop with shared-memory-resident weights (bit-equal outputs required)",
         "# wgrad: three shifted input copies per stage vs one haloed row tile (wide = one N=192 MMA, narrow = 3 x N=64)",
         "# shape [N, H, C, K]        fprop shifted  resident |  fprop+stats shifted  resident |  wgrad copies   wide   narrow | wgrad rel.err wide / narrow"]
for (N, H, C, K) in ((256, 56, 64, 64), (256, 51, 64, 64), (256, 51, 64, 128), (256, 28, 64, 64)):
    x = torch.randn(N, H, H, C, device="cuda").bfloat16()
    w = (torch.randn(K, 9, C, device="cuda") * 0.05).bfloat16()
    dy = torch.randn(N, H, H, K, device="cuda").bfloat16()
    tf, ts, ys = [], [], []
    for mode in (0, 1):
        lib.tdl_conv2d_set_halo_res_mode(mode)
        ys.append(O.conv2d_fprop(x, w, R=3, S=3, pad_h=1, pad_w=1, out_hw=(H, H)))
        tf.append(timed(lambda: O.conv2d_fprop(x, w, R=3, S=3, pad_h=1, pad_w=1, out_hw=(H, H))))
        ts.append(timed(lambda: O.conv2d_fprop_stats(x, w, R=3, S=3, pad_h=1, pad_w=1, out_hw=(H, H))))
    lib.tdl_conv2d_set_halo_res_mode(1)
    same = torch.equal(ys[0], ys[1])
    tw, dws = [], []
    for mode in (0, 1, 2):
        lib.tdl_conv2d_set_wgrad_row_mode(mode)
        dw = torch.zeros(K, 9, C, device="cuda")
        O.conv2d_wgrad(x, dy, dw, R=3, S=3, pad_h=1, pad_w=1)
        dws.append(dw)
        scratch = torch.zeros(K, 9, C, device="cuda")
        tw.append(timed(lambda: O.conv2d_wgrad(x, dy, scratch, R=3, S=3, pad_h=1, pad_w=1)))
    lib.tdl_conv2d_set_wgrad_row_mode(1)
    lines.append(f"  [{N}, {H}, {C}, {K}]".ljust(28) + f"{tf[0]*1e3:9.1f} {tf[1]*1e3:9.1f} |" +
                 f"{ts[0]*1e3:16.1f} {ts[1]*1e3:9.1f} |" + f"{tw[0]*1e3:10.1f} {tw[1]*1e3:8.1f} {tw[2]*1e3:8.1f} |" +
                 f" {rel(dws[1], dws[0]):.2e} / {rel(dws[2], dws[0]):.2e}   fprop bit-equal: {same}")
os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
with open(a.out, "w") as f:
    f.write("\n".join(lines) + "\n")
print("\n".join(lines))
