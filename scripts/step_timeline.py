"""Per-stream busy time of one eager training step (CUPTI kernel records): how much of the weight-gradient / comm side
streams really overlaps the main stream.

    python scripts/step_timeline.py [--model resnet50] [--batch 256] [--size 224]
"""
import argparse
import json
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import ProfilerActivity, profile

from tensorflowdistributedlearning_b200 import models
from tensorflowdistributedlearning_b200.parallel.tower import TowerGroup
from tensorflowdistributedlearning_b200.parallel.trainer import TowerTrainer, classification_loss

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="resnet50")
ap.add_argument("--batch", type=int, default=256)
ap.add_argument("--size", type=int, default=224)
a = ap.parse_args()
group = TowerGroup(device="cuda")
tr = TowerTrainer(models.build(a.model, num_classes=1000), group, batch=a.batch, input_hwc=(a.size, a.size, 3),
                  loss_fn=classification_loss, use_graph=False, kind="sgdm", lr=0.1, momentum=0.9, weight_decay=1e-4)
tr.images.copy_(torch.randint(0, 256, tr.images.shape, dtype=torch.uint8))
tr.labels.copy_(torch.randint(0, 1000, tr.labels.shape))
for _ in range(3):
    tr.run_step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    tr.run_step()
    torch.cuda.synchronize()
path = os.path.join(tempfile.mkdtemp(), "trace.json")
prof.export_chrome_trace(path)
ev = [e for e in json.load(open(path))["traceEvents"] if e.get("cat") == "kernel"]
streams = {}
for e in ev:
    streams.setdefault(e["args"].get("stream"), []).append((e["ts"], e["ts"] + e["dur"], e["name"]))
t0, t1 = min(e["ts"] for e in ev), max(e["ts"] + e["dur"] for e in ev)


def union(iv):
    iv = sorted(iv)
    tot, cs, ce = 0.0, None, None
    for s, e in iv:
        if cs is None:
            cs, ce = s, e
        elif s <= ce:
            ce = max(ce, e)
        else:
            tot += ce - cs
            cs, ce = s, e
    return tot + (ce - cs if cs is not None else 0)


print(f"step span {t1 - t0:.0f} us, {len(ev)} kernels")
for sid, iv in sorted(streams.items(), key=lambda kv: -len(kv[1])):
    busy = union([(s, e) for s, e, _ in iv])
    kinds = {}
    for s, e, n in iv:
        k = n.split("(")[0].split("<")[0][-40:]
        kinds[k] = kinds.get(k, 0) + (e - s)
    top = sorted(kinds.items(), key=lambda kv: -kv[1])[:4]
    print(f"stream {sid}: {len(iv)} kernels, busy {busy:.0f} us; top: " + ", ".join(f"{k} {v:.0f}" for k, v in top))
allbusy = union([(s, e) for iv in streams.values() for s, e, _ in iv])
print(f"any stream busy {allbusy:.0f} us; idle (launch gaps) {t1 - t0 - allbusy:.0f} us")
main = max(streams.values(), key=len)
mainbusy = union([(s, e) for s, e, _ in main])
others = [(s, e) for iv in streams.values() if iv is not main for s, e, _ in iv]
print(f"side-stream work {sum(e - s for s, e in others):.0f} us, of which overlapped with the main stream "
      f"{sum(e - s for s, e in others) + mainbusy - union(others + [(s, e) for s, e, _ in main]):.0f} us")
