"""Per-kernel device-time table for one eager training step (torch.profiler / CUPTI; not an ncu run).

    python scripts/profile_step.py [--model resnet50] [--batch 256] [--out gpurun_out/profile_step.txt]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
from torch.profiler import ProfilerActivity, profile

from tensorflowdistributedlearning_b200 import models
from tensorflowdistributedlearning_b200.parallel.tower import TowerGroup
from tensorflowdistributedlearning_b200.parallel.trainer import TowerTrainer, classification_loss

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="resnet50")
ap.add_argument("--batch", type=int, default=0, help="default: the benchmark batch of the model")
ap.add_argument("--size", type=int, default=0, help="default: the benchmark image size of the model")
ap.add_argument("--out", default="gpurun_out/profile_step.txt")
args = ap.parse_args()
_defaults = {"xception41": (64, 299), "refseg": (32, 101)}  # everything else: 256 x 224x224 (bench.py's configs)
args.batch = args.batch or _defaults.get(args.model, (256, 224))[0]
args.size = args.size or _defaults.get(args.model, (256, 224))[1]

group = TowerGroup(device="cuda")
if args.model == "refseg":  # the reference's own network: 101x101x2 input, Lovasz hinge, Adam
    from tensorflowdistributedlearning_b200.parallel.trainer import segmentation_loss

    model = models.build("refseg", model_name="model")
    tr = TowerTrainer(model, group, batch=args.batch, input_hwc=(101, 101, 2), label_shape=(101, 101, 1),
                      label_dtype=torch.float32, loss_fn=segmentation_loss, use_graph=False, kind="adam", lr=1e-3,
                      mean=[0.47194585] * 2, std=[0.16105755] * 2)
    tr.images.copy_(torch.randint(0, 256, tr.images.shape, dtype=torch.uint8))
    tr.labels.copy_((torch.rand(tr.labels.shape) > 0.6).float())
else:
    model = models.build(args.model, num_classes=1000)
    tr = TowerTrainer(model, group, batch=args.batch, input_hwc=(args.size, args.size, 3),
                      loss_fn=classification_loss, use_graph=False,
                      **(dict(kind="adam", lr=1e-3) if args.model == "xception41" else
                         dict(kind="sgdm", lr=0.1, momentum=0.9, weight_decay=1e-4)))
    tr.images.copy_(torch.randint(0, 256, tr.images.shape, dtype=torch.uint8))
    tr.labels.copy_(torch.randint(0, 1000, tr.labels.shape))
for _ in range(3):
    tr.run_step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(2):
        tr.run_step()
    torch.cuda.synchronize()
rows = []
for e in prof.key_averages():
    t = getattr(e, "self_device_time_total", None)
    if t is None:
        t = getattr(e, "self_cuda_time_total", 0)
    if t:
        rows.append((t / 2.0, e.count // 2, e.key))
rows.sort(reverse=True)
total = sum(r[0] for r in rows)
os.makedirs(os.path.dirname(args.out), exist_ok=True)
with open(args.out, "w") as f:
    f.write(f"# one training step, {args.model} batch {args.batch} {args.size}x{args.size}: total kernel time {total/1e3:.2f} ms\n")
    f.write("# us_per_step  share  launches  kernel\n")
    for t, c, k in rows[:70]:
        f.write(f"{t:10.1f} {100*t/total:6.2f}% {c:5d}  {k[:140]}\n")
print(open(args.out).read())
