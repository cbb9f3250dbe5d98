"""Depthwise -> BatchNorm -> pointwise conv -> BatchNorm (Xception separable conv): every gradient of the folded path
(BatchNorm folded into the pointwise weights) and of the separate-pass path against an fp32 torch reference.

    python scripts/bnfold_check.py
"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from tensorflowdistributedlearning_b200.models.xception import SeparableConv2dSame
from tensorflowdistributedlearning_b200.ops import functional as Fn, layers as L

def rel(a, b): return float((a.float()-b.float()).norm()/b.float().norm().clamp_min(1e-20))
torch.manual_seed(41)
n,h,cin,cout = 4,19,728,728
m = SeparableConv2dSame(cin, cout, 3, 1, 1, scope="s", bn=L.BNConfig(decay=0.9997, eps=1e-3), relu=False, relu_in=True).cuda().train()
for bn in (m.dw_bn, m.pointwise.bn):
    torch.nn.init.uniform_(bn.gamma, 0.5, 1.5); torch.nn.init.normal_(bn.beta, std=0.2)
x0 = torch.randn(n,h,h,cin,device="cuda").bfloat16(); g0 = torch.randn(n,h,h,cout,device="cuda").bfloat16()
res = {}
for flag in (False, True, False):
    Fn.enable_defer_bn(flag); Fn.begin_step()
    for p in m.parameters(): p.grad = None
    st = [b.clone() for b in m.buffers()]
    x = x0.clone().requires_grad_(True)
    y = m(x); y.backward(g0); torch.cuda.synchronize(); Fn.join_side_streams(); torch.cuda.synchronize()
    res[len(res)] = dict(y=y.detach().float(), dx=x.grad.float(), **{k: p.grad.float().clone() for k,p in m.named_parameters()})
    for b,b0 in zip(m.buffers(), st): b.copy_(b0)
# fp32 torch reference
P = {k: p.detach().float().clone().requires_grad_(True) for k,p in m.named_parameters()}
xr = x0.float().clone().requires_grad_(True)
t = F.relu(xr).permute(0,3,1,2)
wd = P["depthwise.weight"].permute(2,0,1).unsqueeze(1)  # [C,1,3,3]
t = F.conv2d(t, wd, None, padding=1, groups=cin)
t = F.batch_norm(t, None, None, P["dw_bn.gamma"], P["dw_bn.beta"], True, 0.1, 1e-3)
wp = P["pointwise.conv.weight"].permute(0,3,1,2)
t = F.conv2d(t, wp)
t = F.batch_norm(t, None, None, P["pointwise.bn.gamma"], P["pointwise.bn.beta"], True, 0.1, 1e-3)
yr = t.permute(0,2,3,1); yr.backward(g0.float())
ref = dict(y=yr.detach(), dx=xr.grad, **{k: v.grad for k,v in P.items()})
for k in ref:
    print(f"{k:28s} nofold {rel(res[0][k], ref[k]):.3e}  fold {rel(res[1][k], ref[k]):.3e}  nofold2 {rel(res[2][k], ref[k]):.3e}  nofold-vs-nofold2 {rel(res[0][k], res[2][k]):.3e}")
