"""Which ATen kernels still run inside a training step, with shapes and the Python frames that launched them.

    python scripts/find_aten_ops.py [--model resnet50] [--batch 256] [--size 224]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
from torch.profiler import ProfilerActivity, profile

from tensorflowdistributedlearning_b200 import models
from tensorflowdistributedlearning_b200.parallel.tower import TowerGroup
from tensorflowdistributedlearning_b200.parallel.trainer import TowerTrainer, classification_loss, segmentation_loss

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="resnet50")
ap.add_argument("--batch", type=int, default=256)
ap.add_argument("--size", type=int, default=224)
a = ap.parse_args()
group = TowerGroup(device="cuda")
if a.model == "refseg":
    tr = TowerTrainer(models.build("refseg", model_name="model"), group, batch=a.batch, input_hwc=(101, 101, 2),
                      label_shape=(101, 101, 1), label_dtype=torch.float32, loss_fn=segmentation_loss, use_graph=False,
                      kind="adam", lr=1e-3, mean=[0.47194585] * 2, std=[0.16105755] * 2)
    tr.labels.copy_((torch.rand(tr.labels.shape) > 0.6).float())
else:
    tr = TowerTrainer(models.build(a.model, num_classes=1000), group, batch=a.batch, input_hwc=(a.size, a.size, 3),
                      loss_fn=classification_loss, use_graph=False, kind="sgdm" if a.model != "xception41" else "adam",
                      lr=0.1 if a.model != "xception41" else 1e-3, momentum=0.9)
    tr.labels.copy_(torch.randint(0, 1000, tr.labels.shape))
tr.images.copy_(torch.randint(0, 256, tr.images.shape, dtype=torch.uint8))
for _ in range(3):
    tr.run_step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU], record_shapes=True, with_stack=True) as prof:
    tr.run_step()
    torch.cuda.synchronize()
agg = {}
for e in prof.events():
    if not e.name.startswith("aten::") or not (e.device_time_total or getattr(e, "cuda_time_total", 0)):
        continue
    t = getattr(e, "self_device_time_total", None) or getattr(e, "self_cuda_time_total", 0)
    if not t:
        continue
    frames = [f for f in (e.stack or []) if "tensorflowdistributedlearning_b200" in f or "bench" in f]
    key = (e.name, str(e.input_shapes)[:80], frames[0][-90:] if frames else "(autograd engine / no python frame)")
    c = agg.setdefault(key, [0, 0.0])
    c[0] += 1
    c[1] += t
for (name, shapes, where), (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{t:9.1f} us  x{n:<3d} {name:28s} {shapes:80s} {where}")
