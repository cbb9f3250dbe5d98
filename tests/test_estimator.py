"""Integration (CPU): Model façade end-to-end, checkpoint layout with TF names, resume, K-fold dirs, export, TTA."""
import json
import os

import numpy as np
import pytest
import torch

from tensorflowdistributedlearning_b200.data import folds as F
from tensorflowdistributedlearning_b200.data import preprocessing as PP
from tensorflowdistributedlearning_b200.estimator import checkpoint as ckpt
from tensorflowdistributedlearning_b200.utils.misc import metric_comparison


def _make_dataset(root, n=24, size=101, seed=0):
    from PIL import Image

    rng = np.random.RandomState(seed)
    os.makedirs(os.path.join(root, "images"), exist_ok=True)
    os.makedirs(os.path.join(root, "masks"), exist_ok=True)
    ids, cov = [], []
    yy, xx = np.mgrid[0:size, 0:size]
    for i in range(n):
        r = rng.uniform(5, 45) if i % 3 else 0
        mask = (((yy - rng.uniform(20, 80)) ** 2 + (xx - rng.uniform(20, 80)) ** 2) < r * r)
        img = np.clip(0.45 + 0.3 * mask + 0.08 * rng.randn(size, size), 0, 1)
        Image.fromarray((img * 255).astype(np.uint8)).save(os.path.join(root, "images", f"id{i:03d}.png"))
        Image.fromarray((mask * 255).astype(np.uint8)).save(os.path.join(root, "masks", f"id{i:03d}.png"))
        ids.append(f"id{i:03d}")
        cov.append(F.coverage_class(mask.astype(np.float32)))
    return np.array(ids, dtype=object), np.array([min(c, 1) for c in cov])


@pytest.fixture(scope="module")
def trained(tmp_path_factory):
    from model import Model

    root = tmp_path_factory.mktemp("salt")
    data_dir = str(root / "data")
    X, y = _make_dataset(data_dir)
    model_dir = str(root / "tinyseg")
    m = Model(model_dir, data_dir, n_gpus=1, n_fold=2, device="cpu", n_blocks=(1, 1, 1), base_depth=16, workers=2,
              save_checkpoints_steps=3, save_summary_steps=2, log_step_count_steps=2, throttle_secs=0)
    res = m.train(X, y, batch_size=4, steps=6)
    return m, model_dir, data_dir, X, y, res


def test_fold_directories_and_symlinks(trained):
    m, model_dir, data_dir, X, y, res = trained
    for mode in ("train", "eval"):
        for kind in ("images", "masks"):
            for k in range(2):
                assert os.path.isdir(os.path.join(model_dir, mode, kind, f"fold{k}"))
    tr = F.list_fold(model_dir, "train", 0)
    ev = F.list_fold(model_dir, "eval", 0)
    assert len(tr) + len(ev) == len(X)
    assert all(os.path.islink(p) and os.path.basename(p) == os.path.basename(q) for p, q in tr)
    assert not ({os.path.basename(p) for p, _ in tr} & {os.path.basename(p) for p, _ in ev})


def test_train_results_and_files(trained):
    m, model_dir, *_rest, res = trained
    assert set(res) == {0, 1}
    for i in (0, 1):
        r = res[i]
        for k in ("metrics/mean_iou", "metrics/mean_acc", "loss/lovasz_loss", "loss", "global_step"):
            assert k in r
        assert r["global_step"] == 6
        fold = os.path.join(model_dir, f"fold{i}")
        assert ckpt.list_checkpoints(fold) == [0, 3, 6]
        assert os.path.exists(os.path.join(fold, "checkpoint"))
        assert os.path.exists(os.path.join(fold, "train", "summaries.jsonl"))
        assert os.path.exists(os.path.join(fold, "eval", "summaries.jsonl"))
        assert os.path.isdir(os.path.join(fold, "export", "best_exporter"))
        # the reference's four image summaries per mode (model.py:406-439): train every save_summary_steps, eval per run
        for mode in ("train", "eval"):
            with open(os.path.join(fold, mode, "summaries.jsonl")) as f:
                recs = [json.loads(line) for line in f]
            tags = {r_["image"] for r_ in recs if "image" in r_}
            assert tags == {f"{mode}_image", f"{mode}_label", f"{mode}_prob", f"{mode}_prediction"}, (mode, tags)
        with open(os.path.join(fold, "train", "summaries.jsonl")) as f:
            steps = sorted({r_["step"] for r_ in map(json.loads, f) if r_.get("image") == "train_prob"})
        assert steps == [2, 4, 6]


def test_checkpoint_has_tf_variable_names_and_layout(trained):
    m, model_dir, *_ = trained
    from safetensors.torch import load_file

    t = load_file(os.path.join(model_dir, "fold0", "model.ckpt-6.safetensors"))
    P = "tinyseg"
    w = t[f"{P}/resnet_v2/conv1_1/weights"]
    assert tuple(w.shape) == (3, 3, 2, 64)  # HWIO
    assert f"{P}/resnet_v2/conv1_1/weights/Adam" in t and f"{P}/resnet_v2/conv1_1/weights/Adam_1" in t
    assert tuple(t[f"{P}/assp/conv/conv_3x3_1_depthwise/depthwise_weights"].shape) == (3, 3, 1024, 1)
    assert int(t["global_step"]) == 6
    assert abs(float(t["beta1_power"]) - 0.9 ** 7) < 1e-6  # TF: beta ** (updates + 1)
    t0 = load_file(os.path.join(model_dir, "fold0", "model.ckpt-0.safetensors"))
    assert 1.0 - float(t0["beta1_power"]) > 0.05 and 1.0 - float(t0["beta2_power"]) > 5e-4  # usable bias correction
    assert f"{P}/resnet_v2/postnorm/moving_variance" in t


def test_resume_equals_uninterrupted(tmp_path):
    """train 4 steps == train 2, 'kill', resume 2 (same data order, CPU determinism)."""
    from tensorflowdistributedlearning_b200 import models
    from tensorflowdistributedlearning_b200.data.dataset import BatchLoader, SyntheticSegmentation
    from tensorflowdistributedlearning_b200.estimator.config import OptimConfig, RunConfig
    from tensorflowdistributedlearning_b200.estimator.estimator import Estimator
    from tensorflowdistributedlearning_b200.parallel.tower import TowerGroup

    group = TowerGroup(device="cpu")
    ds = SyntheticSegmentation(16, input_shape=(33, 33))
    batches = list(zip(range(4), iter(BatchLoader(ds, 4, shuffle=False, pin=False))))
    batches = [b for _, b in batches]

    def builder():
        return models.ResNetSegModel(model_name="m", n_blocks=(1, 1, 1), base_depth=8, input_shape=(33, 33))

    def run(model_dir, chunks):
        for lo, hi in chunks:
            est = Estimator(builder, model_dir, RunConfig(model_dir=model_dir, save_checkpoints_steps=100,
                                                         save_summary_steps=0, log_step_count_steps=0),
                            OptimConfig(lr=1e-3), group=group)
            est.train(lambda: iter(batches[lo:hi]), max_steps=hi)
        return est

    a = run(str(tmp_path / "a"), [(0, 4)])
    b = run(str(tmp_path / "b"), [(0, 2), (2, 4)])
    assert a.global_step == b.global_step == 4
    pa = torch.cat([p.detach().flatten() for p in a.model.parameters()])
    pb = torch.cat([p.detach().flatten() for p in b.model.parameters()])  # padding lanes stay exactly zero in both
    assert torch.allclose(pa, pb, atol=1e-6), float((pa - pb).abs().max())
    # moving statistics, compared on their logical (checkpointed) extent: channel-padding lanes are not state
    from tensorflowdistributedlearning_b200.ops.layers import iter_tf_variables
    sa = torch.cat([t.flatten() for _, t, _ in iter_tf_variables(a.model) if not t.requires_grad])
    sb = torch.cat([t.flatten() for _, t, _ in iter_tf_variables(b.model) if not t.requires_grad])
    assert torch.allclose(sa, sb, atol=1e-6)


def test_predict_tta_average(trained):
    m, model_dir, data_dir, X, *_ = trained
    out = m.predict(os.path.join(data_dir, "images"), batch_size=8, tti=True)
    assert out["probabilities"].shape == (len(X), 101, 101, 1)
    assert len(out["ids"]) == len(X)
    assert set(np.unique(out["mask"])) <= {0.0, 1.0}
    assert 0.0 <= out["probabilities"].min() and out["probabilities"].max() <= 1.0


def test_tta_transformations_are_involutions():
    x = torch.randn(2, 5, 5, 3)
    for t in PP.TRANSFORMATIONS:
        y = PP.single_transformation_from_matrix(x, t)
        assert torch.equal(PP.undo_transformation(y, t), x)
    with pytest.raises(ValueError):
        PP.single_transformation_from_matrix(x, "rot90")


def test_metric_comparison_fixed_and_compat():
    best, cur = {"metrics/mean_iou": 0.5}, {"metrics/mean_iou": 0.6}
    assert metric_comparison(best, cur, "metrics/mean_iou", True) is True
    assert metric_comparison(best, cur, "metrics/mean_iou", True, reference_compat=True) is False  # defect 4
    with pytest.raises(ValueError):
        metric_comparison({}, cur, "metrics/mean_iou")


def test_model_validation_and_params():
    from model import Model

    with pytest.raises(ValueError):
        Model("/tmp/x", "/tmp/y", data_format="NDHWC", device="cpu")


def test_nan_loss_is_fatal_and_summaries_keep_the_reference_tags(tmp_path):
    """a non-finite loss stops the run with the step and tower in the message; the summary hook
    writes the reference's scalar tags (model.py:449-454) to ``<dir>/summaries.jsonl`` every ``save_steps`` steps."""
    import json
    import types

    from tensorflowdistributedlearning_b200.estimator.hooks import NanLossHook, SummarySaverHook

    est = types.SimpleNamespace(group=types.SimpleNamespace(rank=3, is_primary=True))
    NanLossHook().after_step(est, 7, {"loss": 0.25})
    with pytest.raises(FloatingPointError, match="step 7 on tower 3"):
        NanLossHook().after_step(est, 7, {"loss": float("nan")})
    hook = SummarySaverHook(str(tmp_path / "train"), save_steps=20)
    hook.begin(est)
    for step in (19, 20, 40):
        hook.after_step(est, step, {"loss/lovasz_loss": 1.0 / step, "metrics/mean_iou": 0.5, "metrics/mean_acc": 0.9})
    hook.end(est, 40)
    recs = [json.loads(l) for l in open(tmp_path / "train" / "summaries.jsonl")]
    assert [r["step"] for r in recs] == [20, 40]
    assert set(recs[0]) == {"step", "loss/lovasz_loss", "metrics/mean_iou", "metrics/mean_acc"}


def test_tf_tensor_bundle_roundtrip_and_checkpoint_interop(tmp_path):
    """TensorFlow V2 checkpoints (<prefix>.index SSTable + .data shard) read/written without
    TensorFlow; a model saved with ``tf_bundle=True`` loads back from the bundle alone."""
    import struct

    from tensorflowdistributedlearning_b200 import models
    from tensorflowdistributedlearning_b200.estimator import tf_bundle as tb

    assert tb.crc32c(b"123456789") == 0xE3069283 and tb._crc32c_py(b"123456789") == 0xE3069283
    assert tb.unmask_crc(tb.masked_crc(b"abc")) == tb.crc32c(b"abc")
    rng = np.random.default_rng(0)
    tensors = {f"scope/block{i}/unit_{j}/weights": rng.standard_normal((3, 3, 4 + i, 5 + j)).astype(np.float32)
               for i in range(12) for j in range(9)}
    tensors["global_step"] = np.array(1234, dtype=np.int64)
    tensors["beta1_power"] = np.array(0.9 ** 7, dtype=np.float32)
    prefix = str(tmp_path / "model.ckpt-1234")
    tb.write_bundle(prefix, tensors, block_size=512)  # small blocks: several data blocks + a real index block
    raw = open(prefix + ".index", "rb").read()
    assert struct.unpack("<Q", raw[-8:])[0] == 0xDB4775248B80FB57
    header, entries = tb.read_index(prefix)
    assert header["num_shards"] == 1 and header["endianness"] == 0 and len(entries) == len(tensors)
    assert entries["global_step"]["dtype"] == 9 and entries["global_step"]["shape"] == ()
    back = tb.read_bundle(prefix)
    assert set(back) == set(tensors) and all(np.array_equal(back[k], tensors[k]) for k in tensors)
    data_path = prefix + ".data-00000-of-00001"
    blob = bytearray(open(data_path, "rb").read())
    blob[100] ^= 0xFF
    open(data_path, "wb").write(bytes(blob))
    with pytest.raises(IOError, match="checksum"):
        tb.read_bundle(prefix)

    torch.manual_seed(0)
    m1 = models.SmallResNet(num_classes=10)
    d = str(tmp_path / "fold0")
    pfx = ckpt.save_checkpoint(d, m1, None, step=7, tf_bundle=True)
    os.unlink(pfx + ".safetensors")
    assert ckpt.latest_checkpoint(d) == pfx and ckpt.list_checkpoints(d) == [7]
    torch.manual_seed(1)
    m2 = models.SmallResNet(num_classes=10)
    assert ckpt.load_checkpoint(pfx, m2) == 7
    for (n1, p1), (_, p2) in zip(m1.state_dict().items(), m2.state_dict().items()):
        assert torch.equal(p1, p2), n1


def test_cli_eval_command_and_config_overrides(tmp_path, capsys):
    """`eval` evaluates a checkpoint without training (model.py:207-223) and `--set section.field=value` reaches the
    ModelConfig / OptimConfig / RunConfig dataclasses."""
    from tensorflowdistributedlearning_b200.__main__ import main
    from tensorflowdistributedlearning_b200.estimator.config import (ModelConfig, OptimConfig, RunConfig,
                                                                     apply_overrides)

    cfg = {"model": ModelConfig(), "optim": OptimConfig(), "run": RunConfig()}
    apply_overrides(cfg, ["model.base_depth=16", "model.n_blocks=[1,1,1]", "optim.lr=3e-4", "optim.comm=nccl",
                          "run.save_checkpoints_steps=2", "model.input_shape=(37,37)"])
    assert cfg["model"].base_depth == 16 and cfg["model"].n_blocks == (1, 1, 1) and cfg["optim"].lr == 3e-4
    assert cfg["optim"].comm == "nccl" and cfg["run"].save_checkpoints_steps == 2 and cfg["model"].input_shape == (37, 37)
    with pytest.raises(ValueError):
        apply_overrides(cfg, ["model.no_such_field=1"])
    with pytest.raises(ValueError):
        apply_overrides(cfg, ["nosection.lr=1"])

    small = ["--set", "model.base_depth=16", "--set", "model.n_blocks=[1,1,1]", "--set", "model.input_shape=(37,37)",
             "--set", "run.use_cuda_graph=False"]
    d = str(tmp_path / "clirun")
    assert main(["train", "--model", "refseg", "--data", "synthetic", "--model-dir", d, "--steps", "2", "--batch", "2",
                 "--set", "run.save_checkpoints_steps=1", *small]) == 0
    assert ckpt.list_checkpoints(d) == [0, 1, 2]
    capsys.readouterr()
    assert main(["eval", "--model", "refseg", "--data", "synthetic", "--model-dir", d, "--steps", "2", "--batch", "2",
                 *small]) == 0
    out = json.loads(capsys.readouterr().out.strip().splitlines()[-1])
    assert out["global_step"] == 2 and "metrics/mean_iou" in out and "loss/lovasz_loss" in out
    assert main(["eval", "--model", "refseg", "--data", "synthetic", "--model-dir", d, "--steps", "1", "--batch", "2",
                 "--checkpoint", os.path.join(d, "model.ckpt-1"), *small]) == 0
    out1 = json.loads(capsys.readouterr().out.strip().splitlines()[-1])
    assert set(out1) == set(out)
    assert ckpt.list_checkpoints(d) == [0, 1, 2]            # evaluation wrote no checkpoint, trained nothing


def test_model_evaluate_without_training(trained):
    m, model_dir, *_ = trained
    before = {i: ckpt.list_checkpoints(os.path.join(model_dir, f"fold{i}")) for i in (0, 1)}
    res = m.evaluate(batch_size=4)
    assert set(res) == {0, 1}
    for i in (0, 1):
        assert res[i]["global_step"] == 6 and 0.0 <= res[i]["metrics/mean_iou"] <= 1.0
        assert ckpt.list_checkpoints(os.path.join(model_dir, f"fold{i}")) == before[i]
