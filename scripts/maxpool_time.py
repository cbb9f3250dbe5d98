import torch, sys
sys.path.insert(0, ".")
from tensorflowdistributedlearning_b200.ops import pooling
x = torch.randn(256, 112, 112, 64, device="cuda").bfloat16().requires_grad_(True)
def t(fn, it=10):
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it * 1e3
y = pooling.max_pool_same(x, 3, 2)
g = torch.randn_like(y)
print("fwd us", t(lambda: pooling.max_pool_same(x, 3, 2)))
print("bwd us", t(lambda: torch.autograd.grad(y, x, g, retain_graph=True)))
