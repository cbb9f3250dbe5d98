"""GPU end-to-end: the reference-style ``Model`` facade with the native kernels (run with -m gpu)."""
import glob
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_model_facade_trains_checkpoints_and_predicts_on_gpu(tmp_path):
    """model.Model(...).train / predict on synthetic salt-like PNGs: K-fold dirs, GPU input pipeline, CUDA-graph
    trainer, checkpoints with TF names (written as safetensors AND as a TensorFlow V2 bundle), resume, evaluation
    metrics, TTA prediction -- all with the sm_100a library loaded."""
    from model import Model
    from test_estimator import _make_dataset
    from tensorflowdistributedlearning_b200 import _native
    from tensorflowdistributedlearning_b200.estimator import checkpoint as ckpt
    from tensorflowdistributedlearning_b200.estimator import tf_bundle

    before = _native.kernel_launches
    data_dir = str(tmp_path / "data")
    X, y = _make_dataset(data_dir, n=32)
    model_dir = str(tmp_path / "salt")
    m = Model(model_dir, data_dir, n_gpus=1, n_fold=2, n_blocks=(1, 1, 1), base_depth=32, workers=2,
              save_checkpoints_steps=4, save_summary_steps=2, log_step_count_steps=4, throttle_secs=0)
    res = m.train(X, y, batch_size=8, steps=8)
    assert _native.kernel_launches - before > 500, "the native kernels were not used"
    assert set(res) == {0, 1}
    for fold, r in res.items():
        assert np.isfinite(r["loss"]) and 0.0 <= r["metrics/mean_iou"] <= 1.0 and 0.0 <= r["metrics/mean_acc"] <= 1.0
        d = os.path.join(model_dir, f"fold{fold}")
        latest = ckpt.latest_checkpoint(d)
        assert latest is not None and latest.endswith("model.ckpt-8")
        assert glob.glob(os.path.join(d, "train", "*")) and glob.glob(os.path.join(d, "eval", "*"))
    # the same variables as a TensorFlow V2 bundle, read back without TensorFlow
    est = m._estimator(0)
    model = est._ensure_model()
    ckpt.load_checkpoint(ckpt.latest_checkpoint(os.path.join(model_dir, "fold0")), model, None)
    pfx = ckpt.save_checkpoint(str(tmp_path / "tfb"), model, None, step=8, tf_bundle=True)
    bundle = tf_bundle.read_bundle(pfx)
    names = [n for n in bundle if n.endswith("/weights")]
    assert names and all(bundle[n].ndim == 4 for n in names)  # HWIO, logical (un-padded) channel counts
    assert int(bundle["global_step"]) == 8
    pred = m.predict(os.path.join(data_dir, "images"), batch_size=8, tti=True)
    assert len(pred["ids"]) == 32 and pred["probabilities"].shape == (32, 101, 101, 1)
    assert float(pred["probabilities"].min()) >= 0.0 and float(pred["probabilities"].max()) <= 1.0
    assert set(np.unique(pred["mask"])) <= {0.0, 1.0}
    assert torch.cuda.is_available()
