"""Input-pipeline oracles: laplace (1-D compat vs 2-D), projective transform semantics, central crop, augmentation."""
import math

import numpy as np

from tensorflowdistributedlearning_b200.data import preprocessing as PP
from tensorflowdistributedlearning_b200.data.dataset import BatchLoader, SyntheticClassification


def test_laplace_reference_is_1d_along_width():
    x = np.random.RandomState(0).rand(6, 7, 1).astype(np.float32)
    y = PP.laplace(x, reference_compat=True)[:, :, 0]
    a = x[:, :, 0]
    assert np.allclose(y[:, 1:-1], a[:, :-2] - 6 * a[:, 1:-1] + a[:, 2:], atol=1e-6)
    assert np.allclose(y[:, 0], -6 * a[:, 0] + a[:, 1], atol=1e-6)  # zero padding
    y2 = PP.laplace(x, reference_compat=False)[:, :, 0]
    assert not np.allclose(y, y2)
    assert abs(y2[2, 3] - (0.5 * (a[1, 2] + a[1, 4] + a[3, 2] + a[3, 4]) + a[1, 3] + a[3, 3] + a[2, 2] + a[2, 4]
                           - 6 * a[2, 3])) < 1e-5


def test_projective_transform_semantics():
    img = np.arange(25, dtype=np.float64).reshape(5, 5, 1)
    # output (x, y) samples input (x + 1, y): a shift by one pixel, zero outside
    out = PP.transform_image(img, np.array([1, 0, 1, 0, 1, 0, 0, 0.]), "NEAREST")
    assert np.array_equal(out[:, :4, 0], img[:, 1:, 0]) and np.all(out[:, 4, 0] == 0)
    # horizontal flip matrix of the reference: [-1, 0, W, 0, 1, 0, 0, 0] samples x' = W - x (one past the mirror)
    out = PP.transform_image(img, np.array([-1, 0, 5, 0, 1, 0, 0, 0.]), "NEAREST")
    assert np.all(out[:, 0, 0] == 0) and np.array_equal(out[:, 1:, 0], img[:, ::-1, 0][:, :4])
    # rotation by 0 is the identity, by pi maps corners to opposite corners
    t = PP.angles_to_projective_transform(0.0, 5, 5)
    assert np.allclose(PP.transform_image(img, t, "BILINEAR"), img)
    t = PP.angles_to_projective_transform(math.pi, 5, 5)
    assert np.allclose(PP.transform_image(img, t, "BILINEAR")[:, :, 0], img[::-1, ::-1, 0], atol=1e-9)


def test_compose_and_central_crop():
    a = np.array([1, 0, 2, 0, 1, 3, 0, 0.])
    b = np.array([1, 0, -2, 0, 1, -3, 0, 0.])
    assert np.allclose(PP.compose_transforms(a, b), [1, 0, 0, 0, 1, 0, 0, 0])
    assert PP.central_crop(np.zeros((181, 181, 1)), 101 / 181).shape == (101, 101, 1)


def test_read_and_preprocess_shapes_and_determinism():
    rng = np.random.RandomState(1)
    img, mask = rng.rand(101, 101, 1).astype(np.float32), (rng.rand(101, 101, 1) > 0.5).astype(np.float32)
    f0, m0 = PP.read_and_preprocess(img, mask, augment=False)
    assert f0["images"].shape == (101, 101, 2) and m0.shape == (101, 101, 1)
    assert np.allclose(f0["images"][:, :, 0], (img[:, :, 0] - PP.MEAN) / PP.STD, atol=1e-6)
    f1, m1 = PP.read_and_preprocess(img, mask, augment=True, rng=np.random.RandomState(7))
    f2, m2 = PP.read_and_preprocess(img, mask, augment=True, rng=np.random.RandomState(7))
    assert np.array_equal(f1["images"], f2["images"]) and np.array_equal(m1, m2)
    assert f1["images"].shape == (101, 101, 2) and set(np.unique(m1)) <= {0.0, 1.0}


def test_batch_loader_tower_sharding_takes_whole_consecutive_batches():
    ds = SyntheticClassification(n=32, input_shape=(4, 4), num_classes=5)
    all_b = [y.tolist() for _, y in BatchLoader(ds, 4, repeat=False, pin=False)]
    t0 = [y.tolist() for _, y in BatchLoader(ds, 4, repeat=False, rank=0, world=2, pin=False)]
    t1 = [y.tolist() for _, y in BatchLoader(ds, 4, repeat=False, rank=1, world=2, pin=False)]
    assert len(all_b) == 8 and t0 == all_b[0::2] and t1 == all_b[1::2]
