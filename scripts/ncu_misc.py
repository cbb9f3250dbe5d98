"""Driver for ncu captures of single kernels: python scripts/ncu_misc.py <kind>, kind in bn | bn64 | bnwide | lovasz |
optim | adam | stem | stemwgrad | stats1x1 | resmma | conv3x3[_halo] | wgrad64 | pair | depthwise | depthwise_bwd | smallk | maxpool |
upsample | softmaxce | segmetrics | augment | p2p | p2p_wire  (scripts/ncu_all.sh captures one kernel of each)"""
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
elif kind == "wgrad64":
    # weight gradient of ResNet-50's stage-1 3x3 (64 -> 64 @56x56): row mode, the taps of a filter row share one tile
    x = torch.randn(256, 56, 56, 64, device="cuda").bfloat16()
    dy = torch.randn(256, 56, 56, 64, device="cuda").bfloat16()
    dw = torch.zeros(64, 9, 64, device="cuda")
    for _ in range(3):
        O.conv2d_wgrad(x, dy, dw, R=3, S=3, pad_h=1, pad_w=1)
elif kind == "depthwise":
    x = torch.randn(64, 38, 38, 728, device="cuda").bfloat16()
    w = torch.randn(9, 728, device="cuda")
    for _ in range(3):
        O.depthwise3x3_fwd(x, w, None, None, 1, 1, 1, 1, (38, 38), True, False)
elif kind == "depthwise_bwd":
    # Xception middle flow: 3x3 depthwise data gradient (= the forward kernel on flipped taps) and weight gradient
    x = torch.randn(64, 19, 19, 728, device="cuda").bfloat16()
    dy = torch.randn(64, 19, 19, 728, device="cuda").bfloat16()
    w = torch.randn(9, 728, device="cuda")
    for _ in range(3):
        O.depthwise3x3_fwd(dy, w.flip(0).contiguous(), None, None, 1, 1, 1, 1, (19, 19), False, False)
        O.depthwise3x3_wgrad(x, dy, None, 1, 1, 1, 1, True, False)
elif kind == "smallk":
    # the decoder's 3x3 conv to ONE logit channel (core/resnet.py:487-492), batch 32
    x = torch.randn(32, 26, 26, 512, device="cuda").bfloat16()
    w = torch.randn(1, 3, 3, 512, device="cuda") / 68.0
    b = torch.zeros(1, device="cuda")
    dy = torch.randn(32, 26, 26, 1, device="cuda").bfloat16()
    for _ in range(3):
        O.conv2d_smallk_fwd(x, w, b, 1, 1, 1, (26, 26))
        O.conv2d_smallk_bwd_data(dy, w, tuple(x.shape), 1, 1, 1)
        O.conv2d_smallk_bwd_weight(x, dy, torch.zeros_like(w), torch.zeros(1, device="cuda"), 1, 1, 1)
elif kind == "maxpool":
    x = torch.randn(256, 112, 112, 64, device="cuda").bfloat16()
    for _ in range(3):
        y, idx = O.maxpool_fwd(x, 3, 2, 1, 1, (56, 56))
        O.maxpool_bwd(y, idx, tuple(x.shape), 3, 2, 1, 1)
    O.maxpool_fwd_stats(torch.randn(32, 51, 51, 128, device="cuda").bfloat16(), 1, 1, (26, 26))
elif kind == "upsample":
    x = torch.randn(32, 26, 26, 256, device="cuda").bfloat16()
    for _ in range(3):
        y = O.upsample_fwd(x, (101, 101))
        O.upsample_bwd(y, tuple(x.shape), (101, 101))
elif kind == "softmaxce":
    lg = torch.randn(256, 1000, device="cuda")
    lb = torch.randint(0, 1000, (256,), device="cuda")
    for _ in range(3):
        O.softmax_ce(lg, lb, True, 1 / 256, 5)
elif kind == "segmetrics":
    lg = torch.randn(32, 10201, device="cuda")
    lb = (torch.rand(32, 10201, device="cuda") > 0.6).float()
    for _ in range(3):
        O.seg_metrics(lg, lb, 0.0, True)
elif kind == "augment":
    import numpy as np

    from tensorflowdistributedlearning_b200.data.gpu_augment import augment_batch, sample_batch_params

    imgs = torch.randint(0, 256, (256, 101, 101), dtype=torch.uint8, device="cuda")
    msk = torch.randint(0, 2, (256, 101, 101), dtype=torch.uint8, device="cuda") * 255
    params = sample_batch_params(np.random.RandomState(0), 256, 101, 101)
    for _ in range(3):
        augment_batch(imgs, msk, params)
elif kind == "adam":
    n = 41_814_032
    p, gr = torch.randn(n, device="cuda"), torch.randn(n, device="cuda")
    m1, m2 = torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
    st, lr = torch.zeros(1, dtype=torch.int64, device="cuda"), torch.zeros(2, device="cuda")
    for _ in range(3):
        O.lr_schedule_step(st, lr, 1e-3, 0.5, 10000.0)
        O.adam_flat(p, gr, m1, m2, None, lr, 0.9, 0.999, 1e-8, 1.0, True)
elif kind in ("p2p", "p2p_wire"):
    # the fused all-reduce + optimizer kernels with all eight "ranks" on this one GPU (emulation: cross-rank barriers off,
    # every peer pointer is local).  Shows instruction mix / memory behaviour of the reduce + optimizer + all-gather
    # loop; NVLink traffic itself cannot be captured from a single-GPU ncu run.
    from tensorflowdistributedlearning_b200.parallel.p2p import P2PComm, SymmetricBuffer

    world, n = 8, 8 << 20  # 32 MiB fp32 bucket
    gbuf, pbuf = SymmetricBuffer(n * 4, emulate_world=world), SymmetricBuffer(n * 4, emulate_world=world)
    wbuf = SymmetricBuffer(n * 2, emulate_world=world)
    for r in range(world):
        gbuf.tensor(torch.float32, r).normal_()
    comm = P2PComm(gbuf, pbuf, emulate_rank=3)
    comm.set_wire(wbuf)
    m1, m2 = torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
    lr = torch.tensor([0.1, 0.05], device="cuda")
    for _ in range(3):
        if kind == "p2p":
            comm.fused_step(0, n, kind=1, grad_is_bf16=False, zero_grad=True, m1=m1, m2=None, lr_state=lr, a=0.9, b=1e-4,
                            c=0.0, mode=2, max_blocks=64)
            comm.fused_step(0, 1 << 16, kind=2, grad_is_bf16=False, zero_grad=True, m1=m1, m2=m2, lr_state=lr, a=0.9,
                            b=0.999, c=1e-8, mode=1, max_blocks=32)
        else:
            comm.fused_step_wire(0, n, kind=1, zero_grad=True, m1=m1, m2=None, lr_state=lr, a=0.9, b=1e-4, c=0.0, mode=2,
                                 max_blocks=64)
            comm.fused_step_wire(0, 1 << 16, kind=2, zero_grad=True, m1=m1, m2=m2, lr_state=lr, a=0.9, b=0.999, c=1e-8,
                                 mode=1, max_blocks=32)
elif kind == "pair":
    # CTA-pair (cta_group::2) convolution: 3x3 256 -> 256 @14x14, batch 256
    import ctypes

    from tensorflowdistributedlearning_b200 import _native

    _native.lib().tdl_conv2d_set_pair_mode(ctypes.c_int(1))
    x = torch.randn(256, 14, 14, 256, device="cuda").bfloat16()
    w = (torch.randn(256, 9, 256, device="cuda") / 48).bfloat16()
    for _ in range(3):
        O.conv2d_fprop(x, w, R=3, S=3, pad_h=1, pad_w=1, out_hw=(14, 14))
elif kind == "bn64":
    # the narrow, long BatchNorm tensors of ResNet-50's first stage: [802816, 64]
    rows, C = 256 * 56 * 56, 64
    x = torch.randn(rows, C, device="cuda").bfloat16()
    g, b = torch.ones(C, device="cuda"), torch.zeros(C, device="cuda")
    rm, rv = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
    dz = torch.randn(rows, C, device="cuda").bfloat16()
    for _ in range(3):
        mean, invstd, scale, shift = O.bn_fwd_stats(x, g, b, rm, rv, 0.99, 1e-3)
        z = O.bn_apply(x, scale, shift, None, True)
        O.bn_bwd(dz, x, None, scale, shift, mean, invstd, True, False)
elif kind == "bnwide":
    rows, C = 256 * 14 * 14, 1024
    x = torch.randn(rows, C, device="cuda").bfloat16()
    g, b = torch.ones(C, device="cuda"), torch.zeros(C, device="cuda")
    rm, rv = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
    dz = torch.randn(rows, C, device="cuda").bfloat16()
    for _ in range(3):
        mean, invstd, scale, shift = O.bn_fwd_stats(x, g, b, rm, rv, 0.99, 1e-3)
        O.bn_bwd(dz, x, None, scale, shift, mean, invstd, True, False)
torch.cuda.synchronize()
print("ok")
