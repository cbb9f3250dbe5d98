"""Command line: ``python -m tensorflowdistributedlearning_b200 <command>`` (the reference drives everything from
notebooks, Test.ipynb / Untitled.ipynb).

  build                                   compile the sm_100a library
  train   --model refseg --data DIR|synthetic --model-dir OUT --steps N --batch B [--towers T] [--folds K]
  train   --model resnet50 --data synthetic:224x224x3:1000 ...        (classification configs)
  predict --model-dir OUT --data DIR --test-dir DIR [--tta]
  bench   [bench.py arguments]

With ``--towers T`` > 1 the command re-launches itself under torchrun, one process per GPU.
"""
from __future__ import annotations

import argparse
import os
import subprocess
import sys


def _relaunch(towers: int, argv) -> int:
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={towers}", "--master-addr",
           "127.0.0.1", "--master-port", os.environ.get("MASTER_PORT", "29533"), "-m",
           "tensorflowdistributedlearning_b200", *argv]
    return subprocess.call(cmd)


def _ids_and_classes(data_dir: str):
    """Driver recipe without pandas/cv2: ids from the directory listing, coverage class from the masks."""
    import numpy as np

    from .data import folds as F
    from .data import preprocessing as PP

    ids = sorted(os.path.splitext(f)[0] for f in os.listdir(os.path.join(data_dir, "images")) if f.endswith(".png"))
    cls = [F.coverage_class(PP.parse_image(os.path.join(data_dir, "masks", f"{i}.png"))) for i in ids]
    return np.array(ids, dtype=object), np.array(cls)


def cmd_train(a) -> int:
    import torch

    from . import models
    from .data.dataset import BatchLoader, SyntheticClassification, SyntheticSegmentation
    from .estimator.config import OptimConfig, RunConfig
    from .estimator.estimator import Estimator
    from .parallel.tower import TowerGroup

    if a.model == "refseg" and not a.data.startswith("synthetic"):
        from .api import Model

        X, y = _ids_and_classes(a.data)
        m = Model(a.model_dir, a.data, lr=a.lr or 1e-3, n_gpus=a.towers, n_fold=a.folds)
        res = m.train(X, y, a.batch * a.towers, a.steps)
        print(res)
        return 0
    group = TowerGroup()
    if a.model == "refseg":
        ds = SyntheticSegmentation(4096)
        builder = lambda: models.ResNetSegModel(model_name=os.path.basename(os.path.normpath(a.model_dir)))  # noqa: E731
        task, opt, u8 = "segmentation", OptimConfig(kind="adam", lr=a.lr or 1e-3), False
        mean = std = None
    else:
        shape = a.data.split(":")
        h, w, c = (int(v) for v in (shape[1] if len(shape) > 1 else "224x224x3").split("x"))
        classes = int(shape[2]) if len(shape) > 2 else 1000
        import numpy as np

        ds = SyntheticClassification(1 << 16, (h, w), c, classes, dtype=np.uint8)
        builder = lambda: models.build(a.model, num_classes=classes)  # noqa: E731
        task, u8 = "classification", True
        opt = OptimConfig(kind=a.optim, lr=a.lr or (0.1 if a.optim == "sgdm" else 1e-3), decay_steps=0.0,
                          weight_decay=1e-4 if a.optim == "sgdm" else 0.0)
        mean, std = [0.5] * c, [0.25] * c
    est = Estimator(builder, a.model_dir, RunConfig(model_dir=a.model_dir), opt, group=group, task=task,
                    input_is_u8=u8, mean=mean, std=std)
    loader = lambda: iter(BatchLoader(ds, a.batch, shuffle=True, rank=group.rank, world=group.world))  # noqa: E731
    est.train(loader, a.steps)
    if group.is_primary:
        print(f"trained to step {est.global_step}; checkpoints in {a.model_dir}")
    group.shutdown()
    return 0


def cmd_predict(a) -> int:
    import numpy as np

    from .api import Model

    m = Model(a.model_dir, a.data, n_gpus=1, n_fold=a.folds)
    out = m.predict(a.test_dir, a.batch, tti=a.tta)
    np.savez_compressed(os.path.join(a.model_dir, "predictions.npz"), ids=np.array(out["ids"]),
                        probabilities=out["probabilities"], mask=out["mask"])
    print(f"wrote {len(out['ids'])} predictions to {a.model_dir}/predictions.npz")
    return 0


def main(argv=None) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    if argv and argv[0] == "bench":
        bench = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "bench.py")
        return subprocess.call([sys.executable, bench, *argv[1:]])
    ap = argparse.ArgumentParser(prog="tensorflowdistributedlearning_b200")
    sub = ap.add_subparsers(dest="cmd", required=True)
    sub.add_parser("build")
    t = sub.add_parser("train")
    t.add_argument("--model", default="refseg")
    t.add_argument("--data", default="synthetic")
    t.add_argument("--model-dir", required=True)
    t.add_argument("--steps", type=int, default=100)
    t.add_argument("--batch", type=int, default=32, help="per-tower batch")
    t.add_argument("--towers", type=int, default=1)
    t.add_argument("--folds", type=int, default=5)
    t.add_argument("--optim", default="sgdm", choices=["sgdm", "adam"])
    t.add_argument("--lr", type=float, default=0.0)
    p = sub.add_parser("predict")
    p.add_argument("--model-dir", required=True)
    p.add_argument("--data", required=True)
    p.add_argument("--test-dir", required=True)
    p.add_argument("--batch", type=int, default=32)
    p.add_argument("--folds", type=int, default=5)
    p.add_argument("--tta", action="store_true")
    a = ap.parse_args(argv)
    if a.cmd == "build":
        from . import _native

        print(_native.build(force=True))
        return 0
    if a.cmd == "train":
        if a.towers > 1 and "RANK" not in os.environ:
            return _relaunch(a.towers, argv)
        return cmd_train(a)
    return cmd_predict(a)


if __name__ == "__main__":
    sys.exit(main())
