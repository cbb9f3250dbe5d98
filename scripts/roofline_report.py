"""Roofline accounting of one training step from a per-kernel breakdown (scripts/profile_step.py) and the model's own
layer shapes (traced on the CPU with batch 1): ideal FLOPs of the convolution kernels, ideal HBM bytes of the BatchNorm
kernels, against MEASURED_PEAKS.json.  No GPU needed.

    python scripts/roofline_report.py [--model resnet50] [--batch 256] [--size 224] [--breakdown profiles/step_kernel_breakdown_resnet50.txt]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from tensorflowdistributedlearning_b200 import models
from tensorflowdistributedlearning_b200.ops import layers as L

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="resnet50")
ap.add_argument("--batch", type=int, default=256)
ap.add_argument("--size", type=int, default=224)
ap.add_argument("--breakdown", default=None)
a = ap.parse_args()
bd = a.breakdown or f"profiles/step_kernel_breakdown_{a.model}.txt"
peaks = json.load(open("MEASURED_PEAKS.json")) if os.path.exists("MEASURED_PEAKS.json") else {"hbm_gbs": 6550.7, "bf16_tflops": 1632.9}

m = models.build(a.model, num_classes=1000).eval()
conv_macs, bn_elems, dw_macs = 0, 0, 0
hooks = []


def conv_hook(mod, inp, out):
    global conv_macs
    y = out[0] if isinstance(out, tuple) else out
    k, kh, kw, c = mod.weight.shape
    conv_macs += y.numel() // y.shape[-1] * k * kh * kw * c


def dw_hook(mod, inp, out):
    global dw_macs
    y = out[0] if isinstance(out, tuple) else out
    dw_macs += y.numel() * mod.weight.shape[0] * mod.weight.shape[1]


def bn_hook(mod, inp, out):
    global bn_elems
    bn_elems += inp[0].numel()


for mod in m.modules():
    if isinstance(mod, L.Conv2d):
        hooks.append(mod.register_forward_hook(conv_hook))
    elif isinstance(mod, L.DepthwiseConv2d):
        hooks.append(mod.register_forward_hook(dw_hook))
    elif isinstance(mod, L.BatchNorm):
        hooks.append(mod.register_forward_hook(bn_hook))
with torch.no_grad():
    m(torch.zeros(1, a.size, a.size, 3))
B = a.batch
conv_flops_fwd = 2.0 * conv_macs * B
bn_bytes_elem = 2.0  # bf16

fam = {"bn": 0.0, "conv fprop+dgrad": 0.0, "conv wgrad": 0.0, "depthwise": 0.0, "other": 0.0}
for line in open(bd):
    if line.startswith("#"):
        continue
    p = line.split(None, 3)
    if len(p) < 4:
        continue
    us, name = float(p[0]), p[3].strip()
    if name.startswith("_") or name.startswith("aten") or name.startswith("Memcpy"):
        continue
    if "tdl::bn::" in name:
        fam["bn"] += us
    elif "wgrad" in name and "tdl::conv" in name:
        fam["conv wgrad"] += us
    elif "tdl::conv::conv_fprop" in name:
        fam["conv fprop+dgrad"] += us
    elif "tdl::dw" in name:
        fam["depthwise"] += us
    else:
        fam["other"] += us
tot = sum(fam.values())
print(f"# {a.model}, batch {B}, {a.size}x{a.size}: roofline accounting of {bd}")
print(f"# peaks: HBM copy {peaks['hbm_gbs']:.0f} GB/s, cuBLAS bf16 {peaks['bf16_tflops']:.0f} TFLOP/s (burst)")
print(f"kernel time (eager, side streams not overlapped): {tot / 1e3:.2f} ms")
for k, v in fam.items():
    print(f"  {k:18s} {v / 1e3:7.2f} ms  {100 * v / tot:5.1f} %")
t = fam["conv fprop+dgrad"] * 1e-6
if t > 0:
    fl = 2 * conv_flops_fwd  # forward + data gradient
    print(f"conv fprop + dgrad: {fl / 1e12:.2f} TFLOP in {t * 1e3:.2f} ms = {fl / t / 1e12:.0f} TFLOP/s = "
          f"{100 * fl / t / 1e12 / peaks['bf16_tflops']:.0f} % of the cuBLAS burst peak")
t = fam["conv wgrad"] * 1e-6
if t > 0:
    print(f"conv wgrad:         {conv_flops_fwd / 1e12:.2f} TFLOP in {t * 1e3:.2f} ms = {conv_flops_fwd / t / 1e12:.0f} TFLOP/s = "
          f"{100 * conv_flops_fwd / t / 1e12 / peaks['bf16_tflops']:.0f} % of the cuBLAS burst peak")
t = fam["bn"] * 1e-6
if t > 0:
    # forward apply: read x, write z (2 passes); backward reduce: read dz, x (2); backward apply: read dz, x, write dx (3)
    by = 7 * bn_elems * B * bn_bytes_elem
    print(f"BatchNorm:          {by / 1e9:.1f} GB ideal traffic (7 tensor passes over {bn_elems * B / 1e9:.2f} G elements) in "
          f"{t * 1e3:.2f} ms = {by / t / 1e9:.0f} GB/s = {100 * by / t / 1e9 / peaks['hbm_gbs']:.0f} % of copy bandwidth")
if dw_macs:
    print(f"depthwise:          {3 * 2 * dw_macs * B / 1e12:.3f} TFLOP (fwd + dgrad + wgrad) in {fam['depthwise'] / 1e3:.2f} ms")
