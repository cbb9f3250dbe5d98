"""Per-layer check of the fused BatchNorm-backward sums: time of the plain data gradient, of the fused one, and of
the reduction pass (tdl_bn_bwd_reduce) the fusion removes.  Fusion pays when fused - plain < reduce.

    python scripts/bnfuse_check.py
"""
import sys

import torch

sys.path.insert(0, ".")
from tensorflowdistributedlearning_b200.ops import native_ops as O  # noqa: E402

# (name, N, H, C_in of the dgrad (= conv's Cout), K (= conv's Cin = BN channels), R, residual+mask)
SHAPES = [("l1 conv1 dgrad 64->256 tap", 256, 56, 64, 256, 1, True), ("l1 conv2 dgrad 3x3 64->64", 256, 56, 64, 64, 3, False),
          ("l1 conv3 dgrad 256->64", 256, 56, 256, 64, 1, False), ("l2 conv1 dgrad 128->512 tap", 256, 28, 128, 512, 1, True),
          ("l2 conv2 dgrad 3x3 128", 256, 28, 128, 128, 3, False), ("l2 conv3 dgrad 512->128", 256, 28, 512, 128, 1, False),
          ("l3 conv1 dgrad 256->1024 tap", 256, 14, 256, 1024, 1, True), ("l3 conv2 dgrad 3x3 256", 256, 14, 256, 256, 3, False),
          ("l3 conv3 dgrad 1024->256", 256, 14, 1024, 256, 1, False), ("l4 conv1 dgrad 512->2048 tap", 256, 7, 512, 2048, 1, True),
          ("l4 conv2 dgrad 3x3 512", 256, 7, 512, 512, 3, False), ("l4 conv3 dgrad 2048->512", 256, 7, 2048, 512, 1, False)]


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


for name, n, h, c, k, r, tap in SHAPES:
    dy = torch.randn(n, h, h, c, device="cuda").bfloat16()
    w = (torch.randn(k, r * r, c, device="cuda") / (r * r * c) ** 0.5).bfloat16()
    bn_x = torch.randn(n, h, h, k, device="cuda").bfloat16()
    mean, invstd = torch.randn(k, device="cuda"), torch.rand(k, device="cuda") + 0.5
    scale, shift = torch.rand(k, device="cuda") + 0.5, torch.randn(k, device="cuda")
    res = torch.randn(n, h, h, k, device="cuda").bfloat16() if tap else None
    mask = torch.randint(0, 256, (n * h * h, k // 8), dtype=torch.uint8, device="cuda") if tap else None
    pad = (r - 1) // 2
    geo = dict(R=r, S=r, pad_h=pad, pad_w=pad, out_hw=(h, h))
    t_plain = timeit(lambda: O.conv2d_fprop(dy, w, residual=res, out_mask=mask, **geo))
    t_fused = timeit(lambda: O.conv2d_dgrad_bnstats(dy, w, bn_x=bn_x, bn_mean=mean, bn_invstd=invstd, residual=res,
                                                    out_mask=mask, bn_scale=None if tap else scale,
                                                    bn_shift=None if tap else shift, **geo))
    dz2, x2 = bn_x.view(-1, k), bn_x.view(-1, k)
    dgb = torch.empty(2, k, device="cuda")
    t_red = timeit(lambda: O._bn_bwd_reduce(dz2, x2, None, 0, scale, shift, mean, invstd, x2.shape[0], k, not tap, dgb,
                                            None, None))
    print(f"{name:32s} plain {t_plain:7.1f}  fused {t_fused:7.1f}  reduce {t_red:7.1f}  net {t_fused - t_plain - t_red:+7.1f} us",
          flush=True)
