"""Driver for ncu captures of single kernels: python scripts/ncu_misc.py bn|lovasz|optim|stem|stemwgrad|stats1x1|depthwise"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from tensorflowdistributedlearning_b200.ops import native_ops as O

kind = sys.argv[1]
torch.manual_seed(0)
if kind == "bn":
    rows, C = 256 * 56 * 56, 256
    x = torch.randn(rows, C, device="cuda").bfloat16()
    g, b = torch.ones(C, device="cuda"), torch.zeros(C, device="cuda")
    rm, rv = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
    dz = torch.randn(rows, C, device="cuda").bfloat16()
    for _ in range(3):
        mean, invstd, scale, shift = O.bn_fwd_stats(x, g, b, rm, rv, 0.99, 1e-3)
        z = O.bn_apply(x, scale, shift, None, True)
        O.bn_bwd(dz, x, None, scale, shift, mean, invstd, True, False)
elif kind == "lovasz":
    lg = torch.randn(32, 10201, device="cuda")
    lb = (torch.rand(32, 10201, device="cuda") > 0.6).float()
    for _ in range(3):
        O.lovasz_hinge(lg, lb, True, 1 / 32)
elif kind == "optim":
    n = 25_557_032
    p, gr, m = torch.randn(n, device="cuda"), torch.randn(n, device="cuda"), torch.zeros(n, device="cuda")
    st, lr = torch.zeros(1, dtype=torch.int64, device="cuda"), torch.zeros(2, device="cuda")
    for _ in range(3):
        O.lr_schedule_step(st, lr, 0.1)
        O.sgd_momentum_flat(p, gr, m, None, lr, 0.9, 1e-4, 1.0, True)
elif kind == "stem":
    from tensorflowdistributedlearning_b200.ops import functional as Fn

    x = torch.randn(256, 224, 224, 3, device="cuda").bfloat16()
    w = torch.randn(64, 7, 7, 3, device="cuda") / 12.0
    planes = O.conv_stem_pack(x)
    wp = O.conv_stem_weight_prep(w, 3, 3)
    for _ in range(3):
        O.conv_stem_fprop(planes, wp, (224, 224), 64, 7, 7, 3, 3, (112, 112), True)
elif kind == "stemwgrad":
    x = torch.randn(256, 224, 224, 3, device="cuda").bfloat16()
    dy = torch.randn(256, 112, 112, 64, device="cuda").bfloat16()
    planes = O.conv_stem_pack(x)
    for _ in range(3):
        O.conv_stem_wgrad(planes, dy, (224, 224), 3, 7, 7, 3, 3)
elif kind == "bnfuse":
    # data gradient 64 -> 256 @56x56 with residual + mask AND the fused BatchNorm-backward sums (kBwdStats)
    dy = torch.randn(256, 56, 56, 64, device="cuda").bfloat16()
    wd = (torch.randn(256, 1, 64, device="cuda") / 8).bfloat16()
    res = torch.randn(256, 56, 56, 256, device="cuda").bfloat16()
    bx = torch.randn(256, 56, 56, 256, device="cuda").bfloat16()
    mean, invstd = torch.randn(256, device="cuda"), torch.rand(256, device="cuda") + 0.5
    mask = torch.randint(0, 256, (256 * 56 * 56, 32), dtype=torch.uint8, device="cuda")
    for _ in range(3):
        O.conv2d_dgrad_bnstats(dy, wd, R=1, S=1, out_hw=(56, 56), bn_x=bx, bn_mean=mean, bn_invstd=invstd,
                               residual=res, out_mask=mask)
elif kind == "stats1x1":
    x = torch.randn(256, 56, 56, 64, device="cuda").bfloat16()
    w = (torch.randn(256, 1, 64, device="cuda") / 8).bfloat16()
    for _ in range(3):
        O.conv2d_fprop_stats(x, w, R=1, S=1, out_hw=(56, 56))
        O.conv2d_fprop(x, w, R=1, S=1, out_hw=(56, 56))
elif kind == "resmma":
    # data gradient of a unit's first 1x1 conv (64 -> 256 @56x56) with the identity gradient added on the tensor core
    # and the previous unit's ReLU mask applied in the epilogue
    dy = torch.randn(256, 56, 56, 64, device="cuda").bfloat16()
    wd = (torch.randn(256, 1, 64, device="cuda") / 8).bfloat16()
    res = torch.randn(256, 56, 56, 256, device="cuda").bfloat16()
    mask = torch.randint(0, 256, (256 * 56 * 56, 32), dtype=torch.uint8, device="cuda")
    for _ in range(3):
        O.conv2d_fprop(dy, wd, R=1, S=1, out_hw=(56, 56), residual=res, out_mask=mask)
elif kind in ("conv3x3", "conv3x3_halo"):
    from tensorflowdistributedlearning_b200 import _native

    _native.lib().tdl_conv2d_set_halo_mode(1 if kind.endswith("halo") else 0)
    x = torch.randn(256, 56, 56, 64, device="cuda").bfloat16()
    w = (torch.randn(64, 9, 64, device="cuda") / 24).bfloat16()
    for _ in range(3):
        O.conv2d_fprop(x, w, R=3, S=3, pad_h=1, pad_w=1, out_hw=(56, 56))
elif kind == "depthwise":
    x = torch.randn(64, 38, 38, 728, device="cuda").bfloat16()
    w = torch.randn(9, 728, device="cuda")
    for _ in range(3):
        O.depthwise3x3_fwd(x, w, None, None, 1, 1, 1, 1, (38, 38), True, False)
torch.cuda.synchronize()
print("ok")
