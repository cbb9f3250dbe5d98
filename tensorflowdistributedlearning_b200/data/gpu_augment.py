"""Batch augmentation on the GPU (``csrc/augment.cu``): the decoded images cross PCIe once, everything else --
reflect pad, transpose, composed projective warp, crop, normalisation, laplace channel -- happens on the device.

``sample_batch_params`` draws the per-sample parameters on the host with the same generator as the CPU path
(``preprocessing.sample_augmentation``), so both paths produce the same augmentation for the same seed.
"""
from __future__ import annotations

import ctypes
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch

from .. import _native as N
from . import preprocessing as PP


def sample_batch_params(rng: np.random.RandomState, batch: int, height: int, width: int, pad: int = 40, **aug_kw) -> List[Dict]:
    return [PP.sample_augmentation(rng, height + 2 * pad, width + 2 * pad, **aug_kw) for _ in range(batch)]


def augment_batch(images: torch.Tensor, masks: Optional[torch.Tensor], params: List[Dict], pad: int = 40,
                  reference_compat: bool = True) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """images/masks: [B, H, W] (or [B, H, W, 1]) CUDA tensors, decoded and un-normalised: fp32 in [0, 1] or raw uint8
    (divided by 255 inside the kernel, so only one byte per pixel crosses PCIe).  ``params``: per-sample dicts from
    ``preprocessing.sample_augmentation`` (``identity_augmentation()`` for evaluation batches).
    Returns ([B,H,W,2] normalised image + laplace channel, [B,H,W,1] mask), one kernel launch."""
    B, H, W = images.shape[0], images.shape[1], images.shape[2]
    is_u8 = images.dtype == torch.uint8
    want = torch.uint8 if is_u8 else torch.float32
    images = images.reshape(B, H, W).contiguous().to(want)
    if masks is not None:
        masks = masks.reshape(B, H, W).contiguous().to(want)
    dev = images.device
    t = torch.tensor(np.stack([p["transform"] for p in params]).astype(np.float32)).to(dev, non_blocking=True)
    tr = torch.tensor([int(p["transpose"]) for p in params], dtype=torch.int32).to(dev, non_blocking=True)
    br = None
    if any(p.get("brightness", 0.0) for p in params):
        br = torch.tensor([float(p.get("brightness", 0.0)) for p in params], dtype=torch.float32).to(dev, non_blocking=True)
    out = torch.empty((B, H, W, 2), dtype=torch.float32, device=dev)
    out_mask = torch.empty((B, H, W, 1), dtype=torch.float32, device=dev) if masks is not None else None
    rc = N.lib().tdl_augment_seg_ex(N.ptr(images), N.ptr(masks), ctypes.c_int(int(is_u8)), N.ptr(t), N.ptr(tr),
                                    N.ptr(br), N.ptr(out), N.ptr(out_mask), ctypes.c_int(B), ctypes.c_int(H),
                                    ctypes.c_int(W), ctypes.c_int(pad), ctypes.c_float(PP.MEAN), ctypes.c_float(PP.STD),
                                    ctypes.c_int(0 if reference_compat else 1), N.stream_ptr())
    N.check(rc, "tdl_augment_seg_ex")
    return out, out_mask
