"""BatchNorm kernel timings: separate-finalize path vs the cluster-folded path, CUDA events, L2 flushed between
iterations.  Bytes = algorithmic tensor traffic; GB/s against MEASURED_PEAKS.json's copy bandwidth.

    python scripts/bn_bench.py [--out gpurun_out/bn_bench.txt]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from tensorflowdistributedlearning_b200.ops import native_ops as O

ap = argparse.ArgumentParser()
ap.add_argument("--out", default="gpurun_out/bn_bench.txt")
a = ap.parse_args()
peak = 6550.0
try:
    peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                       "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:  # noqa: BLE001
    pass
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timed(fn, iters=10):
    for _ in range(3):
        fn()
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


lines = [f"# BatchNorm kernels, us (median of 10, L2 flushed); GB/s of algorithmic bytes; peak copy bandwidth {peak:.0f} GB/s",
         f"# TDL_BN_CLUSTER={os.environ.get('TDL_BN_CLUSTER', '8')}",
         "# rows C res | apply(old) fused-apply | bwd: partial(old) partial(new) apply(old) fused-apply | frac of peak (new)"]
for rows, C, res in ((802816, 64, False), (802816, 256, True), (200704, 128, False), (200704, 512, True),
                     (50176, 256, False), (50176, 1024, True), (12544, 512, False), (12544, 2048, True)):
    x = (torch.randn(rows, C, device="cuda") * 2 + 0.5).bfloat16()
    r = torch.randn(rows, C, device="cuda").bfloat16() if res else None
    dz = torch.randn(rows, C, device="cuda").bfloat16()
    gamma, beta = torch.rand(C, device="cuda") + 0.5, torch.zeros(C, device="cuda")
    rm, rv = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
    mean, invstd, scale, shift = O.bn_fwd_stats(x, gamma, beta, rm, rv, 0.99, 1e-3)
    partials, nparts = O.bn_partial_stats(x)
    partials = partials[:nparts * 2 * C].clone().view(nparts, 2, C)
    if os.environ.get("BN_BENCH_ROWS"):  # timing experiment: how fast is the streaming part with a short fold
        nparts = min(nparts, int(os.environ["BN_BENCH_ROWS"]))
        partials = partials[:nparts].contiguous()
    z, zmask = O.bn_apply(x, scale, shift, r, True, want_mask=True) if res else (O.bn_apply(x, scale, shift, r, True), None)
    zarg = zmask if res else None
    t_apply_old = timed(lambda: O.bn_apply(x, scale, shift, r, True, want_mask=res))
    t_apply_new = timed(lambda: O.bn_apply_fused(x, partials, nparts, rows, gamma, beta, rm, rv, 0.99, 1e-3, r, True,
                                                 want_mask=res))
    import ctypes
    from tensorflowdistributedlearning_b200 import _native as N
    dgb = torch.empty(2, C, device="cuda")
    scratch = O._bn_scratch(rows, C, x.device)

    def old_partial():
        O._bn_bwd_reduce(dz, x, zarg, int(zarg is not None), scale, shift, mean, invstd, rows, C, True, dgb, None, None)

    npb = ctypes.c_int(0)

    def new_partial():
        N.check(N.lib().tdl_bn_bwd_partial(N.ptr(dz), N.ptr(x), N.ptr(zarg), ctypes.c_int(int(zarg is not None)),
                                           N.ptr(scale), N.ptr(shift), N.ptr(mean), N.ptr(invstd),
                                           ctypes.c_int64(rows), ctypes.c_int(C), ctypes.c_int(1), N.ptr(scratch),
                                           ctypes.c_int64(scratch.numel()), ctypes.byref(npb), N.stream_ptr()),
                "tdl_bn_bwd_partial")

    new_partial()
    bp = scratch[:npb.value * 2 * C].clone().view(npb.value, 2, C)
    if os.environ.get("BN_BENCH_ROWS"):
        npb.value = min(npb.value, int(os.environ["BN_BENCH_ROWS"]))
        bp = bp[:npb.value].contiguous()
    dx = torch.empty_like(x)

    def old_apply():
        N.check(N.lib().tdl_bn_bwd_apply(N.ptr(dz), N.ptr(x), N.ptr(zarg), ctypes.c_int(int(zarg is not None)),
                                         N.ptr(scale), N.ptr(shift), N.ptr(mean), N.ptr(invstd), N.ptr(dgb[0]),
                                         N.ptr(dgb[1]), N.ptr(dx), None, ctypes.c_int64(rows), ctypes.c_int(C),
                                         ctypes.c_int(1), N.stream_ptr()), "tdl_bn_bwd_apply")

    def new_apply():
        N.check(N.lib().tdl_bn_bwd_apply_fused(N.ptr(dz), N.ptr(x), N.ptr(zarg), ctypes.c_int(int(zarg is not None)),
                                               N.ptr(scale), N.ptr(shift), N.ptr(mean), N.ptr(invstd), N.ptr(bp),
                                               ctypes.c_int(npb.value), N.ptr(dgb[0]), N.ptr(dgb[1]), None, None,
                                               N.ptr(dx), None, ctypes.c_int64(rows), ctypes.c_int(C), ctypes.c_int(1),
                                               N.stream_ptr()), "tdl_bn_bwd_apply_fused")

    T = rows * C * 2
    tp_old, tp_new, ta_old, ta_new = timed(old_partial), timed(new_partial), timed(old_apply), timed(new_apply)
    b_apply = T * (3 if res else 2) + (T // 16 if res else 0)
    b_part = 2 * T + (T // 16 if res else 0)
    b_bapply = 3 * T + (T // 16 if res else 0)
    lines.append(f"{rows:7d} {C:5d} {int(res)} | {t_apply_old:7.1f} {t_apply_new:7.1f} | {tp_old:7.1f} {tp_new:7.1f} "
                 f"{ta_old:7.1f} {ta_new:7.1f} | apply {b_apply / t_apply_new / 1e3 / peak:.2f} "
                 f"partial {b_part / tp_new / 1e3 / peak:.2f} bwd-apply {b_bapply / ta_new / 1e3 / peak:.2f}  "
                 f"(old: {b_apply / t_apply_old / 1e3 / peak:.2f} {b_part / tp_old / 1e3 / peak:.2f} "
                 f"{b_bapply / ta_old / 1e3 / peak:.2f})  bwd rows {npb.value}")
os.makedirs(os.path.dirname(a.out), exist_ok=True)
open(a.out, "w").write("\n".join(lines) + "\n")
print("\n".join(lines))
