"""Worker for tests/test_towers.py: runs under torchrun with the gloo backend (CPU towers)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from tensorflowdistributedlearning_b200 import models  # noqa: E402
from tensorflowdistributedlearning_b200.ops import functional as TF  # noqa: E402
from tensorflowdistributedlearning_b200.parallel.optimizer import DistributedOptimizer  # noqa: E402
from tensorflowdistributedlearning_b200.parallel.tower import TowerGroup  # noqa: E402


def main():
    out_dir, mode = sys.argv[1], sys.argv[2]
    group = TowerGroup(device="cpu")
    if mode == "trainer_scalars":
        return trainer_scalars(out_dir, group)
    if mode == "fd_exchange":
        return fd_exchange(out_dir, group)
    torch.manual_seed(100 + group.rank)  # different init per tower: the X2 broadcast must make them equal
    model = models.SmallResNet(num_classes=10)
    if mode == "frozen_bn":
        model.eval()
    else:
        model.train()
    opt = DistributedOptimizer(model, group, kind="adam", lr=1e-2, decay_steps=10000.0, decay_rate=0.5)
    g = torch.Generator().manual_seed(7)
    X = torch.randn(16, 32, 32, 3, generator=g)
    Y = torch.randint(0, 10, (16,), generator=g)
    per = 16 // group.world
    xs, ys = X[group.rank * per:(group.rank + 1) * per], Y[group.rank * per:(group.rank + 1) * per]
    hashes = []
    for step in range(3):
        if mode == "rank1_dies" and group.rank == 1 and step == 1:
            os._exit(17)  # fault injection: a tower disappears mid-run without any goodbye
        opt.begin_step()
        loss, _, _ = TF.softmax_cross_entropy(model(xs), ys)
        loss.backward()
        opt.step()
        hashes.append(opt.flat.param_flat.clone())
    torch.save({"params": opt.flat.param_flat.clone(), "stats": opt.flat.stats_flat.clone(), "per_step": hashes,
                "step": int(opt.global_step.item()), "loss": float(loss)}, os.path.join(out_dir, f"{mode}_w{group.world}_r{group.rank}.pt"))
    group.barrier()
    group.shutdown()


def fd_exchange(out_dir, group):
    """The descriptor hand-over of the VMM symmetric heap (parallel/p2p.exchange_fds) with ordinary files standing in
    for cuMem allocations: all-gather, then a broadcast from tower 0."""
    from tensorflowdistributedlearning_b200.parallel.p2p import exchange_fds

    path = os.path.join(out_dir, f"fdx_{group.rank}.txt")
    with open(path, "w") as f:
        f.write(f"tower {group.rank}")
    fd = os.open(path, os.O_RDONLY)
    got = exchange_fds(fd, group.rank, group.world, None)
    texts = []
    for r in range(group.world):
        # pread: a descriptor received through SCM_RIGHTS shares its file offset with the sender's and the other towers'
        texts.append(os.pread(got[r], 64, 0).decode())
        if r != group.rank:
            os.close(got[r])
    one = exchange_fds(fd if group.rank == 0 else None, group.rank, group.world, None, root=0)
    root_text = os.pread(one[0], 64, 0).decode()
    others_empty = all(one[r] is None for r in range(1, group.world))
    if group.rank != 0:
        os.close(one[0])
    os.close(fd)
    torch.save({"texts": texts, "root_text": root_text, "others_empty": others_empty},
               os.path.join(out_dir, f"fd_exchange_w{group.world}_r{group.rank}.pt"))
    group.barrier()
    group.shutdown()


def trainer_scalars(out_dir, group):
    """X4/X5: the logged loss is the tower mean and the metric sums are global, both out of the step's packed
    reduction (no extra collective)."""
    from tensorflowdistributedlearning_b200.parallel.trainer import TowerTrainer, classification_loss

    torch.manual_seed(5)
    model = models.SmallResNet(num_classes=10)
    tr = TowerTrainer(model, group, batch=8, input_hwc=(32, 32, 3), loss_fn=classification_loss, input_is_u8=False,
                      use_graph=False, kind="sgdm", lr=0.05)
    g = torch.Generator().manual_seed(11 + group.rank)  # every tower sees different data
    rec = []
    for _ in range(2):
        x = torch.randn(8, 32, 32, 3, generator=g)
        y = torch.randint(0, 10, (8,), generator=g)
        logged = tr.train_step(x, y)
        rec.append(dict(logged=logged, local=float(tr.local_loss), top1_mean_of_sums=float(tr.metric_sums["top1"]),
                        top1_local_sum=float(tr.metrics["top1"].float().sum())))
    torch.save(rec, os.path.join(out_dir, f"trainer_scalars_w{group.world}_r{group.rank}.pt"))
    group.barrier()
    group.shutdown()


if __name__ == "__main__":
    main()
