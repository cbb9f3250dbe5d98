"""Driver for ncu captures of the memory-bound kernels: python scripts/ncu_misc.py bn|lovasz|optim"""
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
torch.cuda.synchronize()
print("ok")
