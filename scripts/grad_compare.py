"""Per-parameter gradient comparison, native kernels vs the torch-op path: python scripts/grad_compare.py refseg|xception41|resnet50"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from tensorflowdistributedlearning_b200 import models
from tensorflowdistributedlearning_b200.ops import functional as Fn

which = sys.argv[1] if len(sys.argv) > 1 else "refseg"
torch.manual_seed(4)
if which == "refseg":
    model = models.build("refseg", model_name="model").cuda().train()
    x = torch.randn(8, 101, 101, 2, device="cuda").bfloat16()
    labels = (torch.rand(8, 101, 101, 1, device="cuda") > 0.6).float()
    loss_of = lambda o: Fn.lovasz_loss(labels, o)
else:
    model = models.build(which, num_classes=10).cuda().train()
    if hasattr(model, "keep_prob"):
        model.keep_prob = 1.0
    x = torch.randn(8, 128, 128, 3, device="cuda").bfloat16()
    y = torch.randint(0, 10, (8,), device="cuda")
    loss_of = lambda o: Fn.softmax_cross_entropy(o.float(), y)[0]
for mod in model.modules():
    if getattr(mod, "gamma", None) is not None:
        torch.nn.init.uniform_(mod.gamma, 0.5, 1.5)
out = {}
# three runs on identical weights: native bf16 kernels, torch ops in bf16, torch ops in fp32 (the yardstick for how much
# of any disagreement is just bf16 rounding amplified by batch statistics)
for name, flag, inp in (("native", False, x), ("torch_bf16", True, x), ("torch_fp32", True, x.float())):
    Fn.force_reference(flag)
    Fn.begin_step()
    model.zero_grad(set_to_none=True)
    for mod in model.modules():
        if hasattr(mod, "moving_mean"):
            mod.moving_mean.zero_(); mod.moving_variance.fill_(1.0)
    loss = loss_of(model(inp))
    loss.backward()
    out[name] = (float(loss.detach()), {n: p.grad.float().clone() for n, p in model.named_parameters() if p.grad is not None})
Fn.force_reference(False)
print("loss", {k: round(v[0], 5) for k, v in out.items()})
cosf = lambda a, b: float(torch.nn.functional.cosine_similarity(a.flatten(), b.flatten(), dim=0))
print("  native~fp32  torchbf16~fp32  native~torchbf16   parameter")
names = [n for n, g in out["native"][1].items() if g.dim() >= 3]
for n in names[:6] + names[len(names) // 2 - 2:len(names) // 2 + 2] + names[-6:]:
    g, t, f = out["native"][1][n], out["torch_bf16"][1][n], out["torch_fp32"][1][n]
    print(f"  {cosf(g, f):10.3f}  {cosf(t, f):14.3f}  {cosf(g, t):16.3f}   {n} {tuple(g.shape)}")
allc = lambda a, b: cosf(torch.cat([out[a][1][n].flatten() for n in names]), torch.cat([out[b][1][n].flatten() for n in names]))
print("all conv weights: native~fp32 %.3f  torch_bf16~fp32 %.3f  native~torch_bf16 %.3f" % (
    allc("native", "torch_fp32"), allc("torch_bf16", "torch_fp32"), allc("native", "torch_bf16")))
