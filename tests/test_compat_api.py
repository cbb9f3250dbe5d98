"""The reference's import paths AND call signatures: functional layers with scopes, streaming metrics
returning (value, update_op), ``ignore`` in the Lovasz hinge, ``brightness_range`` in the augmentation, and the
device-augmentation plumbing of the input pipeline (with the CPU oracle injected in place of the CUDA kernel)."""
import numpy as np
import pytest
import torch

from tensorflowdistributedlearning_b200 import compat
from tensorflowdistributedlearning_b200.ops import reference as R


@pytest.fixture(autouse=True)
def _fresh_stores():
    compat.reset_variables()
    compat.reset_local_variables()
    yield
    compat.reset_variables()
    compat.reset_local_variables()


def test_miou_signature_and_streaming_semantics():
    from core.metric import IOU_THRESHOLDS, mIOU, mean_accuracy

    assert len(IOU_THRESHOLDS) == 10
    g = torch.Generator().manual_seed(0)
    y_true = (torch.rand(4, 9, 9, 1, generator=g) > 0.5).float()
    y_pred = (torch.rand(4, 9, 9, 1, generator=g) > 0.5).float()
    coll_v, coll_u = [], []
    value, update = mIOU(y_true, y_pred, metrics_collections=coll_v, updates_collections=coll_u, name="mean_iou_metric")
    want, _ = R.seg_metrics(y_pred.reshape(4, -1) * 2 - 1, y_true.reshape(4, -1), 0.5, True)
    assert value == 0.0 and abs(update - float(want.mean())) < 1e-6     # value = before, update_op = after (A22)
    assert coll_v == [value] and coll_u == [update]
    value2, update2 = mIOU(y_true, y_true, name="mean_iou_metric")     # perfect batch: per-image score 1
    assert abs(value2 - update) < 1e-12
    assert abs(update2 - (float(want.sum()) + 4.0) / 8.0) < 1e-6
    acc_v, acc_u = mean_accuracy(y_true, y_pred, name="mean_acc_metric")
    assert acc_v == 0.0 and abs(acc_u - float((y_true == y_pred).float().mean())) < 1e-6
    compat.reset_local_variables("mean_iou_metric")
    assert mIOU(y_true, y_true, name="mean_iou_metric")[1] == 1.0


def test_split_separable_conv2d_is_a_function_with_scoped_variables():
    from core.layers import _fixed_padding, _upsample, split_separable_conv2d

    x = torch.randn(2, 13, 13, 16)
    y = split_separable_conv2d(x, 8, kernel_size=3, rate=2, scope="assp/conv/conv_3x3_1", is_training=True)
    assert y.shape == (2, 13, 13, 8)
    m = compat.variables("assp/conv/conv_3x3_1")
    names = set(m.depthwise.tf_variables()) | set(m.pointwise.conv.tf_variables())
    assert "assp/conv/conv_3x3_1_depthwise/depthwise_weights" in names
    assert "assp/conv/conv_3x3_1_pointwise/weights" in names
    y2 = split_separable_conv2d(x, 8, rate=2, scope="assp/conv/conv_3x3_1", is_training=False)   # reuse, no new vars
    assert compat.variables("assp/conv/conv_3x3_1") is m and y2.shape == y.shape
    # NCHW at the edge
    yc = split_separable_conv2d(x.permute(0, 3, 1, 2), 8, rate=2, scope="assp/conv/conv_3x3_1", data_format="NCHW",
                                is_training=False)
    assert torch.allclose(yc.permute(0, 2, 3, 1), y2, atol=1e-6)
    p = _fixed_padding(torch.zeros(1, 3, 5, 5), 3)                       # reference default: NCHW
    assert p.shape == (1, 3, 7, 7)
    u = _upsample(torch.ones(1, 2, 13, 13), (26, 26))
    assert u.shape == (1, 2, 26, 26) and torch.allclose(u, torch.ones_like(u))


def test_bottleneck_and_resnet_model_callable_like_the_reference():
    from core.resnet import bottleneck, resnet_model

    ep = {}
    x = torch.randn(2, 8, 8, 32)
    y = bottleneck(x, depth=32, depth_bottleneck=8, stride=2, outputs_collections=ep, scope="b/unit_1", is_training=True)
    assert y.shape == (2, 4, 4, 32) and "b/unit_1/bottleneck_v2" in ep
    img = torch.randn(1, 2, 37, 37)
    logits = resnet_model(img, "m", 0.001, 0.99, 0.001, True, "NCHW", False, 8, 32, (37, 37), (1, 1, 1), "bottleneck")
    assert logits.shape == (1, 1, 37, 37)


def test_lovasz_hinge_ignore_matches_masked_flat_loss():
    from core.losses import flatten_binary_scores, lovasz_hinge, lovasz_loss

    g = torch.Generator().manual_seed(3)
    logits = torch.randn(3, 7, 7, generator=g)
    labels = (torch.rand(3, 7, 7, generator=g) > 0.5).float()
    labels[0, :2] = 255.0                                                # void pixels
    got = lovasz_hinge(logits, labels, per_image=True, ignore=255.0)
    want = torch.stack([R.lovasz_hinge_flat(lg.reshape(-1)[lb.reshape(-1) != 255], lb.reshape(-1)[lb.reshape(-1) != 255])
                        for lg, lb in zip(logits, labels)]).mean()
    assert torch.allclose(got, want)
    s, l = flatten_binary_scores(logits, labels, ignore=255.0)
    assert s.numel() == l.numel() == 3 * 49 - 14
    flat = lovasz_hinge(logits, labels, per_image=False, ignore=255.0)
    assert torch.allclose(flat, R.lovasz_hinge_flat(s, l))
    clean = (labels != 255).float() * labels
    a = lovasz_loss(clean.unsqueeze(1), logits.unsqueeze(1), data_format="NCHW")
    b = lovasz_loss(clean.unsqueeze(-1), logits.unsqueeze(-1), data_format="NHWC")
    assert torch.allclose(a, b) and torch.allclose(a, R.lovasz_loss(clean, logits))


def test_brightness_range_shifts_the_normalised_image_before_the_warp():
    from preprocessing.preprocessing import read_and_preprocess

    rng = np.random.RandomState(0)
    img = rng.rand(101, 101, 1).astype(np.float32)
    mask = (rng.rand(101, 101, 1) > 0.5).astype(np.float32)
    kw = dict(horizontal_flip=False, vertical_flip=False, rotate_range=0, height_shift_range=0, width_shift_range=0,
              crop_probability=0.0)
    f0, m0 = read_and_preprocess(img, mask, augment=True, rng=np.random.RandomState(5), brightness_range=0.0, **kw)
    f1, m1 = read_and_preprocess(img, mask, augment=True, rng=np.random.RandomState(5), brightness_range=0.3, **kw)
    d = f1["images"][:, :, 0] - f0["images"][:, :, 0]
    assert np.allclose(d, d.flat[0], atol=1e-5) and 0 < abs(d.flat[0]) <= 0.3   # one constant delta per sample
    assert np.array_equal(m0, m1)


def test_device_augmentation_pipeline_reproduces_the_cpu_pipeline(tmp_path):
    """BatchLoader(device_augment=True): workers decode uint8 only and the batch is augmented in one call with the
    parameters the CPU path would have drawn.  The CPU oracle stands in for the CUDA kernel here (the kernel itself is
    checked against the same oracle in tests/test_gpu_kernels.py)."""
    from PIL import Image

    from tensorflowdistributedlearning_b200.data import preprocessing as PP
    from tensorflowdistributedlearning_b200.data.dataset import BatchLoader, SegmentationFolder

    rng = np.random.RandomState(1)
    pairs = []
    for i in range(6):
        a = (rng.rand(101, 101) * 255).astype(np.uint8)
        b = ((rng.rand(101, 101) > 0.5) * 255).astype(np.uint8)
        Image.fromarray(a).save(tmp_path / f"i{i}.png")
        Image.fromarray(b).save(tmp_path / f"m{i}.png")
        pairs.append((str(tmp_path / f"i{i}.png"), str(tmp_path / f"m{i}.png")))
    calls = []

    def oracle(images_u8, masks_u8, params, reference_compat=True):
        calls.append(len(params))
        assert images_u8.dtype == torch.uint8 and masks_u8.dtype == torch.uint8
        outs = [PP.apply_augmentation(im.numpy()[:, :, None].astype(np.float32) / 255.0,
                                      mk.numpy()[:, :, None].astype(np.float32) / 255.0, p, reference_compat)
                for im, mk, p in zip(images_u8, masks_u8, params)]
        return (torch.from_numpy(np.stack([o[0] for o in outs])), torch.from_numpy(np.stack([o[1] for o in outs])))

    for augment in (True, False):
        ds = SegmentationFolder(pairs, augment=augment, seed=3)
        host = list(BatchLoader(ds, 3, repeat=False, workers=2, pin=False))
        dev = list(BatchLoader(ds, 3, repeat=False, workers=2, pin=False, device_augment=True, augment_fn=oracle,
                               device=torch.device("cpu")))
        assert len(host) == len(dev) == 2
        for (hx, hy), (dx, dy) in zip(host, dev):
            assert torch.allclose(hx, dx, atol=1e-5) and torch.equal(hy, dy)
    assert calls == [3, 3, 3, 3]                                          # ONE augmentation call per batch


def test_single_pass_loader_drops_the_ragged_round():
    """5 batches over 2 towers: both towers must see 2 batches (a third on tower 0 would be a step its peer never
    joins)."""
    from tensorflowdistributedlearning_b200.data.dataset import BatchLoader, SyntheticClassification

    ds = SyntheticClassification(20, (4, 4), 3, 10)
    n = [len(list(BatchLoader(ds, 4, repeat=False, rank=r, world=2, workers=1, pin=False))) for r in (0, 1)]
    assert n == [2, 2]
