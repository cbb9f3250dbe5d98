"""Oracle self-consistency: the CPU restatements of the TF semantics."""
import math

import numpy as np
import pytest
import torch

from tensorflowdistributedlearning_b200.ops import reference as R


def test_same_pad_matches_tf_rules():
    # 224, k=3, s=2 -> 0 before / 1 after (A1)
    assert R.same_pad(224, 3, 2) == (112, 0, 1)
    assert R.same_pad(224, 7, 2) == (112, 2, 3)
    assert R.same_pad(101, 3, 2) == (51, 1, 1)
    assert R.same_pad(13, 3, 1, 8) == (13, 8, 8)
    assert R.same_pad(51, 3, 2) == (26, 1, 1)


def test_max_pool_same_ignores_padding():
    x = -torch.ones(1, 4, 4, 1)
    y = R.max_pool_same_nhwc(x, 3, 2)
    assert y.shape == (1, 2, 2, 1)
    assert torch.all(y == -1)  # zero padding would have produced 0


def test_subsample_starts_at_zero():
    x = torch.arange(25.).view(1, 5, 5, 1)
    y = R.subsample(x, 2)
    assert y.flatten().tolist() == [0, 2, 4, 10, 12, 14, 20, 22, 24]


def test_batch_norm_tf_semantics():
    torch.manual_seed(0)
    x = torch.randn(4, 3, 3, 5) * 3 + 1
    rm, rv = torch.zeros(5), torch.ones(5)
    y = R.batch_norm_train(x, torch.ones(5), torch.zeros(5), rm, rv, 0.99, 1e-3)
    xf = x.reshape(-1, 5)
    assert torch.allclose(rm, 0.01 * xf.mean(0), atol=1e-6)
    assert torch.allclose(rv, 0.99 + 0.01 * xf.var(0, unbiased=True), atol=1e-6)
    assert torch.allclose(y.reshape(-1, 5).mean(0), torch.zeros(5), atol=1e-5)
    yi = R.batch_norm_infer(x, torch.ones(5), torch.zeros(5), rm, rv, 1e-3)
    assert torch.allclose(yi, (x - rm) / torch.sqrt(rv + 1e-3), atol=1e-5)


def test_legacy_bilinear_no_half_pixel():
    x = torch.tensor([0., 1., 2., 3.]).view(1, 1, 4, 1)
    y = R.resize_bilinear_legacy(x, (1, 8)).flatten()
    # src = dst * 0.5 -> 0, .5, 1, 1.5, 2, 2.5, 3, 3 (clamped)
    assert torch.allclose(y, torch.tensor([0, .5, 1, 1.5, 2, 2.5, 3, 3.]))


def test_upsample_reference_shape_and_constant():
    x = torch.full((2, 13, 13, 3), 2.5)
    y = R.upsample_reference(x, (26, 26))
    assert y.shape == (2, 26, 26, 3)
    assert torch.allclose(y, torch.full_like(y, 2.5))


def _lovasz_numpy(logits, labels):
    signs = 2.0 * labels - 1.0
    errors = 1.0 - logits * signs
    order = np.argsort(-errors, kind="stable")
    es, gs = errors[order], labels[order]
    gts = gs.sum()
    inter = gts - np.cumsum(gs)
    union = gts + np.cumsum(1 - gs)
    jac = 1.0 - inter / union
    jac[1:] = jac[1:] - jac[:-1]
    return float(np.dot(np.maximum(es, 0), jac))


def test_lovasz_matches_numpy_and_finite_differences():
    rng = np.random.RandomState(0)
    lg = rng.randn(3, 40)
    lb = (rng.rand(3, 40) > 0.6).astype(np.float64)
    lb[1] = 0
    t = torch.tensor(lg, requires_grad=True)
    loss = R.lovasz_hinge_per_image(t, torch.tensor(lb))
    for i in range(3):
        assert abs(float(loss[i]) - _lovasz_numpy(lg[i], lb[i])) < 1e-9
    loss.sum().backward()
    eps = 1e-6
    for (i, j) in [(0, 3), (2, 17), (1, 5)]:
        p = lg.copy(); p[i, j] += eps
        m = lg.copy(); m[i, j] -= eps
        fd = (_lovasz_numpy(p[i], lb[i]) - _lovasz_numpy(m[i], lb[i])) / (2 * eps)
        assert abs(fd - float(t.grad[i, j])) < 1e-5


def test_lovasz_empty():
    assert float(R.lovasz_hinge_flat(torch.zeros(0), torch.zeros(0))) == 0.0


def test_seg_metrics_reference_compat_factor():
    # one image, IoU = 0.75 -> thresholds 0.5..0.7 pass (5 of 10)
    labels = torch.zeros(1, 8); labels[0, :4] = 1
    logits = torch.full((1, 8), -5.0); logits[0, :3] = 5.0
    iou_c, acc = R.seg_metrics(logits, labels, 0.5, True)
    iou_k, _ = R.seg_metrics(logits, labels, 0.5, False)
    assert abs(float(iou_k) - 0.5) < 1e-6
    assert abs(float(iou_c) - 0.5 * 0.75) < 1e-6
    assert abs(float(acc) - 7 / 8) < 1e-6
    # empty prediction on empty mask scores 1
    iou_e, _ = R.seg_metrics(torch.full((1, 8), -5.0), torch.zeros(1, 8), 0.5, False)
    assert float(iou_e) == 1.0


def test_adam_tf_and_decay():
    p = torch.tensor([1.0]); g = torch.tensor([0.5]); m = torch.zeros(1); v = torch.zeros(1)
    R.adam_tf_step(p, g, m, v, 0.001, 1)
    lr_t = 0.001 * math.sqrt(1 - 0.999) / (1 - 0.9)
    expect = 1.0 - lr_t * 0.05 / (math.sqrt(0.001 * 0.25) + 1e-8)
    assert abs(float(p) - expect) < 1e-7
    assert abs(R.exponential_decay(0.001, 10000, 10000, 0.5) - 0.0005) < 1e-12
    assert abs(R.exponential_decay(0.001, 5000, 10000, 0.5) - 0.001 * 0.5 ** 0.5) < 1e-12


def test_variance_scaling_statistics():
    w = torch.empty(256, 3, 3, 128)
    R.variance_scaling_(w)
    std = math.sqrt(1.3 * 2 / (3 * 3 * 128))
    assert float(w.abs().max()) <= 2 * std + 1e-6
    # truncated normal at 2 sigma has std ~0.88 sigma
    assert abs(float(w.std()) / std - 0.8796) < 0.02


def test_topk_ties_lower_index_wins():
    logits = torch.tensor([[1.0, 1.0, 0.0]])
    assert float(R.topk_correct(logits, torch.tensor([0]), 1)) == 1.0
    assert float(R.topk_correct(logits, torch.tensor([1]), 1)) == 0.0
    assert float(R.topk_correct(logits, torch.tensor([1]), 2)) == 1.0
