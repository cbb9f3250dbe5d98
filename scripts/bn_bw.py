"""Achieved bandwidth of the BatchNorm kernels at ResNet-50 shapes (CUDA events, distinct tensors, 10 iterations).

    python scripts/bn_bw.py
"""
import sys

import torch

sys.path.insert(0, ".")
from tensorflowdistributedlearning_b200.ops import native_ops as O  # noqa: E402

SHAPES = [(802816, 64), (802816, 256), (200704, 128), (200704, 512), (50176, 256), (50176, 1024), (12544, 512), (12544, 2048)]


def timeit(fn, iters=10):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


print("rows x C            stats us (GB/s)   apply us (GB/s)   bwd-reduce us (GB/s)   bwd-apply us (GB/s)")
for rows, C in SHAPES:
    x = torch.randn(rows, C, device="cuda").bfloat16()
    dz = torch.randn(rows, C, device="cuda").bfloat16()
    g, b = torch.ones(C, device="cuda"), torch.zeros(C, device="cuda")
    rm, rv = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
    mean, invstd, scale, shift = O.bn_fwd_stats(x, g, b, rm, rv, 0.99, 1e-3)
    nb = rows * C * 2
    t_stats = timeit(lambda: O.bn_fwd_stats(x, g, b, rm, rv, 0.99, 1e-3))
    t_apply = timeit(lambda: O.bn_apply(x, scale, shift, None, True))
    dgb = torch.empty(2, C, device="cuda")
    t_red = timeit(lambda: O._bn_bwd_reduce(dz, x, None, 0, scale, shift, mean, invstd, rows, C, True, dgb, None, None))
    t_all = timeit(lambda: O.bn_bwd(dz, x, None, scale, shift, mean, invstd, True, False))
    t_bapply = t_all - t_red
    gb = lambda passes, us: passes * nb / us / 1e3  # noqa: E731
    print(f"{rows:7d} x {C:4d}   {t_stats:7.1f} ({gb(1, t_stats):5.0f})   {t_apply:7.1f} ({gb(2, t_apply):5.0f})   "
          f"{t_red:7.1f} ({gb(2, t_red):5.0f})        {t_bapply:7.1f} ({gb(3, t_bapply):5.0f})", flush=True)
