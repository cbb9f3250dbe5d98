"""CUDA-event time of the Lovasz hinge kernel at the reference shape (32 images of 101x101)."""
import sys

import torch

sys.path.insert(0, ".")
from tensorflowdistributedlearning_b200.ops import native_ops as O  # noqa: E402

for B, P in ((32, 10201), (32, 4096), (256, 10201)):
    lg = torch.randn(B, P, device="cuda")
    lb = (torch.rand(B, P, device="cuda") > 0.6).float()
    for _ in range(3):
        O.lovasz_hinge(lg, lb, True, 1 / B)
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(20):
        O.lovasz_hinge(lg, lb, True, 1 / B)
    e1.record()
    torch.cuda.synchronize()
    print(f"B={B} P={P}: {e0.elapsed_time(e1) / 20 * 1e3:.1f} us per call")
