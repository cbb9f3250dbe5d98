"""Command line: ``python -m tensorflowdistributedlearning_b200 <command>`` (the reference drives everything from
notebooks, Test.ipynb / Untitled.ipynb).

  build                                   compile the sm_100a library
  train   --model refseg --data DIR|synthetic --model-dir OUT --steps N --batch B [--towers T] [--folds K]
  train   --model resnet50 --data synthetic:224x224x3:1000 ...        (classification configs)
  eval    --model refseg --data DIR|synthetic --model-dir OUT [--checkpoint PREFIX] [--steps N]
                                          evaluate a checkpoint without training (model.py:207-223)
  predict --model-dir OUT --data DIR --test-dir DIR [--tta]
  bench   [bench.py arguments]

``train`` / ``eval`` accept ``--set section.field=value`` (repeatable) for any field of the ModelConfig / OptimConfig /
RunConfig dataclasses, e.g. ``--set model.base_depth=128 --set optim.lr=3e-4 --set run.save_checkpoints_steps=100``.

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


def _configs(a):
    """ModelConfig / OptimConfig / RunConfig for a train / eval command line, `--set` overrides applied."""
    from .estimator.config import ModelConfig, OptimConfig, RunConfig, apply_overrides

    seg = a.model == "refseg"
    shape = a.data.split(":")
    classes = None if seg else (int(shape[2]) if len(shape) > 2 else 1000)
    optim = getattr(a, "optim", "sgdm")
    lr = getattr(a, "lr", 0.0)
    model = ModelConfig(name=a.model, num_classes=classes)
    opt = (OptimConfig(kind="adam", lr=lr or 1e-3) if seg else
           OptimConfig(kind=optim, lr=lr or (0.1 if optim == "sgdm" else 1e-3), decay_steps=0.0,
                       weight_decay=1e-4 if optim == "sgdm" else 0.0))
    run = RunConfig(model_dir=a.model_dir)
    apply_overrides({"model": model, "optim": opt, "run": run}, getattr(a, "set", None))
    return model, opt, run


def _synthetic_estimator(a, group):
    """Estimator + dataset for the synthetic configurations (every BASELINE.json config has one)."""
    import numpy as np

    from .data.dataset import SyntheticClassification, SyntheticSegmentation
    from .estimator.estimator import Estimator

    model, opt, run = _configs(a)
    name = os.path.basename(os.path.normpath(a.model_dir))
    if a.model == "refseg":
        ds = SyntheticSegmentation(4096, tuple(model.input_shape), model.in_channels)
        task, u8, mean, std = "segmentation", False, None, None
    else:
        shape = a.data.split(":")
        h, w, c = (int(v) for v in (shape[1] if len(shape) > 1 else "224x224x3").split("x"))
        ds = SyntheticClassification(1 << 16, (h, w), c, model.num_classes, dtype=np.uint8)
        task, u8, mean, std = "classification", True, [0.5] * c, [0.25] * c
    est = Estimator(lambda: model.build(name), a.model_dir, run, opt, group=group, task=task, input_is_u8=u8,
                    mean=mean, std=std)
    return est, ds


def _model_kwargs(a):
    """`--set` overrides mapped onto the `Model(**kwargs)` facade (real-data segmentation runs)."""
    from dataclasses import asdict

    model, opt, run = _configs(a)
    kw = {k: v for k, v in asdict(model).items() if k not in ("name", "in_channels", "num_classes")}
    kw.update({k: getattr(run, k) for k in ("save_checkpoints_steps", "save_summary_steps", "log_step_count_steps",
                                            "use_cuda_graph")})
    kw.update({k: getattr(opt, k) for k in ("comm", "overlap", "bucket_cap_mb")})
    return kw, opt, run


def cmd_train(a) -> int:
    from .data.dataset import BatchLoader
    from .parallel.tower import TowerGroup

    if a.model == "refseg" and not a.data.startswith("synthetic"):
        from .api import Model

        X, y = _ids_and_classes(a.data)
        kw, opt, run = _model_kwargs(a)
        m = Model(a.model_dir, a.data, lr=opt.lr, n_gpus=a.towers, n_fold=a.folds, seed=run.seed, **kw)
        res = m.train(X, y, a.batch * a.towers, a.steps)
        print(res)
        return 0
    group = TowerGroup()
    est, ds = _synthetic_estimator(a, group)
    loader = lambda: iter(BatchLoader(ds, a.batch, shuffle=True, rank=group.rank, world=group.world))  # noqa: E731
    est.train(loader, a.steps)
    if group.is_primary:
        print(f"trained to step {est.global_step}; checkpoints in {a.model_dir}")
    group.shutdown()
    return 0


def cmd_eval(a) -> int:
    """Evaluate a checkpoint without training: the reference reaches this through tf.estimator's evaluate on the
    EvalSpec input (model.py:207-223).  Prints and returns the metric dict of the newest (or the named) checkpoint."""
    import json

    from .data.dataset import BatchLoader
    from .parallel.tower import TowerGroup

    group = TowerGroup()
    if a.model == "refseg" and not a.data.startswith("synthetic"):
        from .api import Model

        kw, opt, run = _model_kwargs(a)
        m = Model(a.model_dir, a.data, lr=opt.lr, n_gpus=group.world, n_fold=a.folds, seed=run.seed, group=group, **kw)
        results = m.evaluate(a.batch, steps=a.steps or None, checkpoint_path=a.checkpoint)
    else:
        est, ds = _synthetic_estimator(a, group)
        loader = lambda: iter(BatchLoader(ds, a.batch, shuffle=False, repeat=False, rank=group.rank,  # noqa: E731
                                          world=group.world))
        results = est.evaluate(loader, steps=a.steps or max(len(ds) // a.batch, 1), checkpoint_path=a.checkpoint,
                               sharded=True)
    if group.is_primary:
        print(json.dumps(results, default=float))
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
    t.add_argument("--set", action="append", default=[], metavar="SECTION.FIELD=VALUE",
                   help="override a ModelConfig / OptimConfig / RunConfig field (model.*, optim.*, run.*)")
    e = sub.add_parser("eval")
    e.add_argument("--model", default="refseg")
    e.add_argument("--data", default="synthetic")
    e.add_argument("--model-dir", required=True)
    e.add_argument("--checkpoint", default=None, help="checkpoint prefix (default: the newest in --model-dir)")
    e.add_argument("--steps", type=int, default=0, help="evaluation batches (default: one pass)")
    e.add_argument("--batch", type=int, default=32, help="per-tower batch")
    e.add_argument("--towers", type=int, default=1)
    e.add_argument("--folds", type=int, default=5)
    e.add_argument("--set", action="append", default=[], metavar="SECTION.FIELD=VALUE")
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
    if a.cmd == "eval":
        if a.towers > 1 and "RANK" not in os.environ:
            return _relaunch(a.towers, argv)
        return cmd_eval(a)
    return cmd_predict(a)


if __name__ == "__main__":
    sys.exit(main())
