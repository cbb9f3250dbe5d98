"""Weight-gradient kernel on the ResNet-50 layer shapes (batch 256): epilogue (per-thread red.global.add vs bulk
reductions), whole-wave grid snapping, and the split-count model constant.  CUDA events, L2 flushed, median of 20.

    python scripts/wgrad_bench.py [--out gpurun_out/wgrad_bench.txt]
"""
import argparse
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from tensorflowdistributedlearning_b200 import _native
from tensorflowdistributedlearning_b200.ops import native_ops as O

ap = argparse.ArgumentParser()
ap.add_argument("--out", default="gpurun_out/wgrad_bench.txt")
a = ap.parse_args()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
lib = _native.lib()
lib.tdl_conv2d_set_wgrad_tuning.argtypes = [ctypes.c_int, ctypes.c_float]


def timed(fn, iters=20, warmup=3):
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


shapes = [(256, 56, 64, 64, 1, 1), (256, 56, 64, 64, 3, 1), (256, 56, 64, 256, 1, 1), (256, 56, 256, 64, 1, 1),
          (256, 28, 128, 128, 3, 1), (256, 28, 512, 128, 1, 1), (256, 28, 128, 512, 1, 1), (256, 14, 256, 256, 3, 1),
          (256, 14, 1024, 256, 1, 1), (256, 14, 256, 1024, 1, 1), (256, 7, 512, 512, 3, 1), (256, 7, 2048, 512, 1, 1),
          (256, 7, 512, 2048, 1, 1), (256, 56, 128, 128, 3, 2), (256, 56, 256, 512, 1, 2), (256, 14, 1024, 2048, 1, 2)]
# (bulk, snap, model_k, pair)
configs = [(0, 0, 1.1, 0), (1, 1, 1.1, 0), (1, 1, 1.1, 1), (1, 1, 2.2, 1), (1, 1, 4.4, 1), (1, 1, 0.0, 1)]
lines = ["# wgrad us per call; columns = (epilogue: red/bulk, wave snap, split-model constant; 0 = model off, p1 = CTA-pair kernel for Cout >= 256)",
         "# shape [N,H,C,K,R,stride]".ljust(34) + "".join(f"{('bulk' if b else 'red')}/s{s}/k{k:g}/p{p}".rjust(17) for b, s, k, p in configs) + "   max rel.diff vs first"]
tot = [0.0] * len(configs)
for (N, H, C, K, R, stride) in shapes:
    x = torch.randn(N, H, H, C, device="cuda").bfloat16()
    P, pt, pb = O.same_pad(H, R, stride)
    dy = torch.randn(N, P, P, K, device="cuda").bfloat16()
    row, ref, worst = [], None, 0.0
    for i, (b, s, k, pr) in enumerate(configs):
        lib.tdl_conv2d_set_wgrad_bulk_mode(b)
        lib.tdl_conv2d_set_wgrad_pair_mode(pr)
        lib.tdl_conv2d_set_wgrad_tuning(s, ctypes.c_float(k))
        dw = torch.zeros(K, R * R, C, device="cuda")
        O.conv2d_wgrad(x, dy, dw, R=R, S=R, stride=stride, pad_h=pt, pad_w=pt)
        if ref is None:
            ref = dw
        else:
            worst = max(worst, float((dw - ref).norm() / ref.norm()))
        scratch = torch.zeros(K, R * R, C, device="cuda")
        t = timed(lambda: O.conv2d_wgrad(x, dy, scratch, R=R, S=R, stride=stride, pad_h=pt, pad_w=pt)) * 1e3
        row.append(t)
        tot[i] += t
    lines.append(f"  {[N, H, C, K, R, stride]}".ljust(34) + "".join(f"{t:17.1f}" for t in row) + f"   {worst:.1e}")
lines.append("  sum".ljust(34) + "".join(f"{t:17.1f}" for t in tot))
lib.tdl_conv2d_set_wgrad_bulk_mode(1)
lib.tdl_conv2d_set_wgrad_pair_mode(0)
lib.tdl_conv2d_set_wgrad_tuning(1, ctypes.c_float(-1.0))
os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
open(a.out, "w").write("\n".join(lines) + "\n")
print("\n".join(lines))
