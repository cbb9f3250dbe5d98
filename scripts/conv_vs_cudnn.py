"""ResNet-50 layer shapes at batch 256: this library's fprop / data gradient / weight gradient vs cuDNN (torch,
channels-last bf16, cudnn.benchmark autotuning ON).  CUDA events, L2 flushed between iterations, median of 20.

    python scripts/conv_vs_cudnn.py [--out gpurun_out/conv_layer_timings_vs_cudnn.txt]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from tensorflowdistributedlearning_b200.ops import native_ops as O

ap = argparse.ArgumentParser()
ap.add_argument("--out", default="gpurun_out/conv_layer_timings_vs_cudnn.txt")
a = ap.parse_args()
torch.backends.cudnn.benchmark = True
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


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


shapes = [(256, 56, 64, 64, 1, 1), (256, 56, 64, 64, 3, 1), (256, 56, 64, 256, 1, 1), (256, 56, 256, 64, 1, 1),
          (256, 28, 128, 128, 3, 1), (256, 28, 512, 128, 1, 1), (256, 28, 128, 512, 1, 1), (256, 14, 256, 256, 3, 1),
          (256, 14, 1024, 256, 1, 1), (256, 14, 256, 1024, 1, 1), (256, 7, 512, 512, 3, 1), (256, 7, 2048, 512, 1, 1),
          (256, 7, 512, 2048, 1, 1), (256, 56, 128, 128, 3, 2), (256, 56, 256, 512, 1, 2)]
lines = ["# ResNet-50 layer shapes at batch 256, CUDA-event timed, L2 flushed between iterations (256 MB write), median of 20",
         "# shape = [N, H=W, Cin, Cout, R=S, stride]; 'cudnn' = torch conv2d / convolution_backward, channels-last bf16,",
         "# torch.backends.cudnn.benchmark = True (autotuned library baseline); TFLOP/s is this library's fprop",
         "# shape                          fprop ms   cudnn ms  TFLOP/s |   dgrad ms   cudnn ms |   wgrad ms   cudnn ms"]
for (N, H, C, K, R, stride) in shapes:
    x = torch.randn(N, H, H, C, device="cuda").bfloat16()
    w = (torch.randn(K, R * R, C, device="cuda") * 0.05).bfloat16()
    P, pt, pb = O.same_pad(H, R, stride)
    pad = (R - 1) // 2
    t_f = timed(lambda: O.conv2d_fprop(x, w, R=R, S=R, stride=stride, pad_h=pad, pad_w=pad, out_hw=(P, P)))
    xc, wc = x.permute(0, 3, 1, 2), w.view(K, R, R, C).permute(0, 3, 1, 2)
    t_fc = timed(lambda: F.conv2d(xc, wc, None, stride=stride, padding=pad))
    dy = torch.randn(N, P, P, K, device="cuda").bfloat16()
    wd = w.float().view(K, R * R, C).flip(1).permute(2, 1, 0).contiguous().bfloat16()
    if stride == 1:
        t_d = timed(lambda: O.conv2d_fprop(dy, wd, R=R, S=R, pad_h=R - 1 - pad, pad_w=R - 1 - pad, out_hw=(H, H)))
    else:
        t_d = timed(lambda: O.conv2d_dgrad_strided(dy, wd, (N, H, H, C), R=R, S=R, stride=stride, pad_h=pad, pad_w=pad))
    dyc = dy.permute(0, 3, 1, 2)
    t_dc = timed(lambda: torch.ops.aten.convolution_backward(dyc, xc, wc, None, [stride, stride], [pad, pad], [1, 1],
                                                             False, [0, 0], 1, [True, False, False]))
    dw = torch.zeros(K, R * R, C, device="cuda")
    t_w = timed(lambda: O.conv2d_wgrad(x, dy, dw, R=R, S=R, stride=stride, pad_h=pad, pad_w=pad))
    t_wc = timed(lambda: torch.ops.aten.convolution_backward(dyc, xc, wc, None, [stride, stride], [pad, pad], [1, 1],
                                                             False, [0, 0], 1, [False, True, False]))
    fl = 2.0 * N * P * P * K * C * R * R
    lines.append(f"  {str([N, H, C, K, R, stride]):30s} {t_f:9.3f} {t_fc:10.3f} {fl / t_f / 1e9:8.0f} | {t_d:10.3f} {t_dc:10.3f} |"
                 f" {t_w:10.3f} {t_wc:10.3f}")
    print(lines[-1], flush=True)
os.makedirs(os.path.dirname(a.out), exist_ok=True)
open(a.out, "w").write("\n".join(lines) + "\n")
