"""Tiny driver for ncu captures of single conv kernels: python scripts/ncu_conv.py fprop|wgrad N H C K R stride [iters]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from tensorflowdistributedlearning_b200.ops import native_ops as O

kind = sys.argv[1]
N, H, C, K, R, stride = map(int, sys.argv[2:8])
iters = int(sys.argv[8]) if len(sys.argv) > 8 else 3
torch.manual_seed(0)
x = torch.randn(N, H, H, C, device="cuda").bfloat16()
w = (torch.randn(K, R * R, C, device="cuda") * 0.05).bfloat16()
pad = (R - 1) // 2
P = (H + 2 * pad - R) // stride + 1
dy = torch.randn(N, P, P, K, device="cuda").bfloat16()
dw = torch.zeros(K, R * R, C, device="cuda")
for _ in range(iters):
    if kind == "fprop":
        O.conv2d_fprop(x, w, R=R, S=R, stride=stride, pad_h=pad, pad_w=pad, out_hw=(P, P))
    else:
        O.conv2d_wgrad(x, dy, dw, R=R, S=R, stride=stride, pad_h=pad, pad_w=pad)
torch.cuda.synchronize()
print("ok")
