"""Image decode / normalise / augmentation / TTA transforms (CPU path; also the oracle for the CUDA warp).

Re-design of the reference's ``preprocessing/preprocessing.py``:

* ``MEAN`` / ``STD``                         -- :7-8
* ``laplace``                                -- :11-30 (a 1-D [1,-6,1] filter along W, defect 8; true 2-D behind a flag)
* ``parse_image`` / ``read_image``           -- :91-101
* ``single_transformation*``                 -- :104-109, :249-278 (TTA: vertical / horizontal / transpose / none)
* ``read_and_preprocess``                    -- :112-246 (reflect-pad 40, transpose, flips, rotation, shift,
                                                optional crop-zoom, projective warp, central crop, laplace channel)
* projective helpers (TF contrib.image)      -- SURVEY appendix A11-A14

Arrays are HWC float32 numpy / torch CPU tensors; batching and device placement live in ``data/dataset.py``.
The shift augmentation is drawn per sample (the reference draws it once per run at graph-build time,
defect 9); ``reference_compat=True`` restores the constant shift.
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Tuple

import numpy as np
import torch

MEAN = 0.47194585
STD = 0.16105755


# ---------------------------------------------------------------------------------------------
# decode / normalise / laplace
# ---------------------------------------------------------------------------------------------
def parse_image(filename: str, size: Tuple[int, int] = (101, 101)) -> np.ndarray:
    """Read a PNG/JPEG as single-channel float32 in [0,1], shape [H, W, 1] (reference ``_parse_image``)."""
    from PIL import Image

    with Image.open(filename) as im:
        arr = np.asarray(im.convert("L"), dtype=np.float32) / 255.0
    if arr.shape != tuple(size):
        raise ValueError(f"{filename}: expected {size}, got {arr.shape}")
    return arr[:, :, None]


def parse_image_u8(filename: str, size: Tuple[int, int] = (101, 101)) -> np.ndarray:
    """The decoded single-channel image as raw uint8 [H, W] (what the device-side augmentation ingests)."""
    from PIL import Image

    with Image.open(filename) as im:
        arr = np.asarray(im.convert("L"), dtype=np.uint8)
    if arr.shape != tuple(size):
        raise ValueError(f"{filename}: expected {size}, got {arr.shape}")
    return arr


def normalise(image: np.ndarray) -> np.ndarray:
    return (image - MEAN) / STD


def laplace(x: np.ndarray, reference_compat: bool = True) -> np.ndarray:
    """x: [H, W, 1].  reference_compat: what the reference's code computes -- [1,-6,1] along W with zero padding
    (its depthwise_conv2d sees [batch=H, height=W, width=1]); otherwise the intended 2-D kernel
    [[.5,1,.5],[1,-6,1],[.5,1,.5]] with zero ('SAME') padding."""
    a = x[:, :, 0]
    if reference_compat:
        p = np.pad(a, ((0, 0), (1, 1)))
        out = p[:, :-2] - 6.0 * p[:, 1:-1] + p[:, 2:]
    else:
        k = np.array([[0.5, 1.0, 0.5], [1.0, -6.0, 1.0], [0.5, 1.0, 0.5]], dtype=a.dtype)
        p = np.pad(a, 1)
        out = np.zeros_like(a)
        for r in range(3):
            for s in range(3):
                out += k[r, s] * p[r:r + a.shape[0], s:s + a.shape[1]]
    return out[:, :, None].astype(x.dtype)


def read_image(image_file: str, mask_file: str):
    return parse_image(image_file), parse_image(mask_file)


# ---------------------------------------------------------------------------------------------
# projective transforms (tf.contrib.image semantics)
# ---------------------------------------------------------------------------------------------
def angles_to_projective_transform(theta: float, height: int, width: int) -> np.ndarray:
    """A12: rotation about the image centre as a flat 8-vector mapping output -> input coordinates."""
    c, s = math.cos(theta), math.sin(theta)
    xo = ((width - 1) - (c * (width - 1) - s * (height - 1))) / 2.0
    yo = ((height - 1) - (s * (width - 1) + c * (height - 1))) / 2.0
    return np.array([c, -s, xo, s, c, yo, 0.0, 0.0], dtype=np.float64)


def flat_to_matrix(t: np.ndarray) -> np.ndarray:
    return np.append(np.asarray(t, dtype=np.float64), 1.0).reshape(3, 3)


def matrix_to_flat(m: np.ndarray) -> np.ndarray:
    m = m / m[2, 2]
    return m.reshape(-1)[:8]


def compose_transforms(*transforms) -> np.ndarray:
    """A13: product M1 . M2 ... Mn of the 3x3 forms, returned flat."""
    m = np.eye(3)
    for t in transforms:
        m = m @ flat_to_matrix(t)
    return matrix_to_flat(m)


def transform_image(img: np.ndarray, t: np.ndarray, interpolation: str = "BILINEAR") -> np.ndarray:
    """A11: output pixel (x, y) samples input ((a0 x + a1 y + a2)/k, (b0 x + b1 y + b2)/k), k = c0 x + c1 y + 1;
    samples outside the image contribute 0.  img: [H, W, C]."""
    h, w, c = img.shape
    ys, xs = np.meshgrid(np.arange(h, dtype=np.float64), np.arange(w, dtype=np.float64), indexing="ij")
    k = t[6] * xs + t[7] * ys + 1.0
    sx = (t[0] * xs + t[1] * ys + t[2]) / k
    sy = (t[3] * xs + t[4] * ys + t[5]) / k
    out = np.zeros_like(img)
    if interpolation == "NEAREST":
        xi = np.rint(sx).astype(np.int64)
        yi = np.rint(sy).astype(np.int64)
        ok = (xi >= 0) & (xi < w) & (yi >= 0) & (yi < h)
        out[ok] = img[yi[ok], xi[ok]]
        return out
    x0 = np.floor(sx).astype(np.int64)
    y0 = np.floor(sy).astype(np.int64)
    fx = (sx - x0)[..., None]
    fy = (sy - y0)[..., None]

    def tap(yy, xx):
        ok = (xx >= 0) & (xx < w) & (yy >= 0) & (yy < h)
        v = np.zeros((h, w, c), dtype=np.float64)
        v[ok] = img[yy[ok], xx[ok]]
        return v

    top = tap(y0, x0) * (1 - fx) + tap(y0, x0 + 1) * fx
    bot = tap(y0 + 1, x0) * (1 - fx) + tap(y0 + 1, x0 + 1) * fx
    return (top * (1 - fy) + bot * fy).astype(img.dtype)


def central_crop(img: np.ndarray, fraction: float) -> np.ndarray:
    """A14 (newer TF1 formula): start = int((dim - dim*f)/2), size = dim - 2*start."""
    h, w = img.shape[:2]
    sh = int((h - h * fraction) / 2)
    sw = int((w - w * fraction) / 2)
    return img[sh:h - sh, sw:w - sw]


# ---------------------------------------------------------------------------------------------
# augmentation
# ---------------------------------------------------------------------------------------------
def sample_augmentation(rng: np.random.RandomState, height: int, width: int, horizontal_flip=True, vertical_flip=True,
                        rotate_range=10, crop_probability=0.0, crop_min_percent=0.9, crop_max_percent=1.1,
                        height_shift_range=0.2, width_shift_range=0.2, constant_shift=None,
                        brightness_range=0.0) -> Dict:
    """Draw one sample's augmentation parameters and the composed output->input transform.
    ``brightness_range`` > 0 adds a delta ~ U(-range, range) to the normalised image before the warp
    (tf.image.random_brightness, preprocessing/preprocessing.py:169-170; the reference's default 0 draws nothing)."""
    ts = []
    do_transpose = bool(rng.rand() < 0.5)
    brightness = float(rng.uniform(-brightness_range, brightness_range)) if brightness_range > 0 else 0.0
    if horizontal_flip and rng.rand() < 0.5:
        ts.append(np.array([-1, 0, width, 0, 1, 0, 0, 0], dtype=np.float64))
    if vertical_flip and rng.rand() < 0.5:
        ts.append(np.array([1, 0, 0, 0, -1, height, 0, 0], dtype=np.float64))
    if rotate_range:
        ang = math.radians(rng.uniform(-rotate_range, rotate_range))
        ts.append(angles_to_projective_transform(ang, height, width))
    if height_shift_range or width_shift_range:
        if constant_shift is not None:
            tx, ty = constant_shift
        else:
            # the reference scales BOTH shifts by `height` (preprocessing.py:196-203); kept
            tx = rng.uniform(-width_shift_range, width_shift_range) * height
            ty = rng.uniform(-height_shift_range, height_shift_range) * height
        ts.append(np.array([1, 0, tx, 0, 1, ty, 0, 0], dtype=np.float64))
    if crop_probability > 0 and rng.rand() < crop_probability:
        pct = rng.uniform(crop_min_percent, crop_max_percent)
        left = rng.uniform(0, width * (1 - pct))
        top = rng.uniform(0, height * (1 - pct))
        ts.append(np.array([pct, 0, left, 0, pct, top, 0, 0], dtype=np.float64))
    t = compose_transforms(*ts) if ts else np.array([1, 0, 0, 0, 1, 0, 0, 0], dtype=np.float64)
    return {"transpose": do_transpose, "transform": t, "brightness": brightness}


def apply_augmentation(image: np.ndarray, mask: np.ndarray, params: Dict, reference_compat: bool = True, pad: int = 40):
    """Deterministic part of the augmentation: normalise -> reflect pad -> (transpose) -> warp -> central crop ->
    laplace channel, for given ``params`` = {'transpose': bool, 'transform': 8-vector}.  Oracle of the CUDA kernel."""
    image = normalise(image)
    h0 = image.shape[0]
    image = np.pad(image, ((pad, pad), (pad, pad), (0, 0)), mode="reflect")
    mask = np.pad(mask, ((pad, pad), (pad, pad), (0, 0)), mode="reflect")
    h = image.shape[0]
    if params["transpose"]:
        image, mask = image.transpose(1, 0, 2), mask.transpose(1, 0, 2)
    if params.get("brightness", 0.0):
        image = image + np.float32(params["brightness"])
    image = transform_image(image, params["transform"], "BILINEAR")
    mask = transform_image(mask, params["transform"], "NEAREST")
    image, mask = central_crop(image, h0 / h), central_crop(mask, h0 / h)
    image = np.concatenate([image, laplace(image, reference_compat)], axis=-1).astype(np.float32)
    return image, mask.astype(np.float32)


def identity_augmentation() -> Dict:
    """Parameters under which the augmentation kernels reduce to normalise + laplace (evaluation batches)."""
    return {"transpose": False, "transform": np.array([1, 0, 0, 0, 1, 0, 0, 0], dtype=np.float64), "brightness": 0.0}


def read_and_preprocess(image: np.ndarray, mask: np.ndarray, augment: bool = False,
                        rng: Optional[np.random.RandomState] = None, reference_compat: bool = True, **aug_kw):
    """image/mask: [H, W, 1] in [0,1].  Returns ({'images': [H,W,2]}, mask[H,W,1]) like the reference."""
    if augment:
        rng = rng or np.random.RandomState()
        pad = 40
        p = sample_augmentation(rng, image.shape[0] + 2 * pad, image.shape[1] + 2 * pad, **aug_kw)
        image, mask = apply_augmentation(image, mask, p, reference_compat, pad)
        return {"images": image}, mask
    image = normalise(image)
    image = np.concatenate([image, laplace(image, reference_compat)], axis=-1).astype(np.float32)
    return {"images": image}, mask.astype(np.float32)


# ---------------------------------------------------------------------------------------------
# test-time augmentation
# ---------------------------------------------------------------------------------------------
TRANSFORMATIONS = ("none", "vertical", "horizontal", "transpose")


def single_transformation_from_matrix(x, transformation: str):
    """x: [..., H, W, C] numpy or torch.  'vertical' = flip up/down, 'horizontal' = flip left/right."""
    is_t = isinstance(x, torch.Tensor)
    if transformation == "vertical":
        return x.flip(-3) if is_t else x[..., ::-1, :, :].copy()
    if transformation == "horizontal":
        return x.flip(-2) if is_t else x[..., :, ::-1, :].copy()
    if transformation == "transpose":
        return x.transpose(-3, -2) if is_t else np.swapaxes(x, -3, -2).copy()
    if transformation == "none":
        return x
    raise ValueError(f"Unknown transformation {transformation}")


def single_transformation(x, transformation: str) -> Dict:
    return {"images": single_transformation_from_matrix(x, transformation)}


def undo_transformation(y, transformation: str):
    """All four TTA transforms are involutions, so undoing == applying again (reference: model.py:379-387)."""
    return single_transformation_from_matrix(y, transformation)


def single_transformation_from_jpeg(filename: str, transformation: str, reference_compat: bool = True) -> Dict:
    img = normalise(parse_image(filename))
    img = np.concatenate([img, laplace(img, reference_compat)], axis=-1).astype(np.float32)
    return single_transformation(img, transformation)
