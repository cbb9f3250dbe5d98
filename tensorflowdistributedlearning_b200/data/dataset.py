"""Datasets and the host-side input pipeline (replaces the tf.data pipeline of model.py:257-324).

* ``SegmentationFolder``  -- (image, mask) pairs of one fold, decode + normalise + (optional) augmentation +
  laplace channel per sample (``data/preprocessing.py``); pairing is by id.
* ``SyntheticClassification`` / ``SyntheticSegmentation`` -- deterministic random data of a named shape, used by
  every BASELINE.json config (there is no network for datasets).
* ``BatchLoader(device_augment=True)`` -- the worker threads only decode; one ``tdl_augment_seg_ex`` launch per
  batch does reflect pad / transpose / warp / crop / normalise / laplace on the GPU (``data/gpu_augment.py``), with
  the per-sample parameters drawn exactly as the CPU path draws them (same seed -> same augmentation).
* ``BatchLoader``         -- shuffle buffer (10 x batch, crossing epoch boundaries like ``shuffle_and_repeat``),
  worker threads for decode/augment (the reference maps single-threaded, model.py:315), pinned batches,
  prefetch depth 2 x towers.  Every tower dequeues whole consecutive batches (A17): tower r of N takes
  batches r, r+N, r+2N, ... of the one global stream.
"""
from __future__ import annotations

import queue
import threading
from typing import Callable, Iterator, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import preprocessing as PP


class SegmentationFolder:
    def __init__(self, pairs: Sequence[Tuple[str, str]], augment: bool = False, reference_compat: bool = True,
                 seed: int = 0, input_shape=(101, 101), **aug_kw):
        self.pairs = list(pairs)
        self.augment, self.reference_compat, self.seed = augment, reference_compat, seed
        self.input_shape = tuple(input_shape)
        aug_kw.setdefault("crop_probability", 0.0)  # the reference's caller disables crop-zoom (model.py:317)
        self.aug_kw = aug_kw

    def __len__(self):
        return len(self.pairs)

    def __getitem__(self, i: int, epoch: int = 0):
        img_path, mask_path = self.pairs[i]
        image = PP.parse_image(img_path, self.input_shape)
        mask = PP.parse_image(mask_path, self.input_shape)
        rng = self._rng(i, epoch)
        feats, mask = PP.read_and_preprocess(image, mask, self.augment, rng, self.reference_compat, **self.aug_kw)
        return feats["images"], mask

    def _rng(self, i: int, epoch: int) -> np.random.RandomState:
        return np.random.RandomState((self.seed * 1_000_003 + epoch * 7919 + i) % (2 ** 31 - 1))

    def raw(self, i: int, epoch: int = 0, pad: int = 40):
        """(image uint8 [H,W], mask uint8 [H,W], augmentation parameters) -- the decode-only half of ``__getitem__``;
        the parameters are drawn from the same generator state, so the device path reproduces the CPU path."""
        img_path, mask_path = self.pairs[i]
        image = PP.parse_image_u8(img_path, self.input_shape)
        mask = PP.parse_image_u8(mask_path, self.input_shape)
        if self.augment:
            h, w = self.input_shape
            params = PP.sample_augmentation(self._rng(i, epoch), h + 2 * pad, w + 2 * pad, **self.aug_kw)
        else:
            params = PP.identity_augmentation()
        return image, mask, params


class SyntheticSegmentation:
    """Salt-shaped synthetic data: blobby masks and images correlated with them, 101x101x2 by default."""

    def __init__(self, n: int = 256, input_shape=(101, 101), channels: int = 2, seed: int = 0):
        self.n, self.input_shape, self.channels, self.seed = n, tuple(input_shape), channels, seed

    def __len__(self):
        return self.n

    def __getitem__(self, i: int, epoch: int = 0):
        rng = np.random.RandomState(self.seed * 977 + i)
        h, w = self.input_shape
        yy, xx = np.mgrid[0:h, 0:w]
        cy, cx, r = rng.uniform(0, h), rng.uniform(0, w), rng.uniform(0, 0.6 * h)
        mask = (((yy - cy) ** 2 + (xx - cx) ** 2) < r * r).astype(np.float32)
        if rng.rand() < 0.3:
            mask[:] = 0
        base = 0.5 * mask + 0.25 * rng.randn(h, w).astype(np.float32)
        img = np.stack([base] + [PP.laplace(base[:, :, None])[:, :, 0] for _ in range(self.channels - 1)], -1)
        return img.astype(np.float32), mask[:, :, None]


class SyntheticClassification:
    def __init__(self, n: int = 1024, input_shape=(32, 32), channels: int = 3, num_classes: int = 10, seed: int = 0,
                 dtype=np.float32):
        self.n, self.input_shape, self.channels, self.num_classes, self.seed = n, tuple(input_shape), channels, num_classes, seed
        self.dtype = dtype

    def __len__(self):
        return self.n

    def __getitem__(self, i: int, epoch: int = 0):
        rng = np.random.RandomState(self.seed * 977 + i)
        label = int(rng.randint(self.num_classes))
        h, w = self.input_shape
        x = rng.randn(h, w, self.channels).astype(np.float32) * 0.5
        x[:, :, label % self.channels] += 0.5 + 0.1 * label  # a learnable signal
        if self.dtype == np.uint8:
            x = np.clip((x * 0.25 + 0.5) * 255.0, 0, 255).astype(np.uint8)
        return x, np.int64(label)


def _collate(samples):
    xs = torch.from_numpy(np.stack([s[0] for s in samples]))
    ys = torch.from_numpy(np.stack([np.asarray(s[1]) for s in samples]))
    return xs, ys


class BatchLoader:
    """Infinite (``repeat``) or single-pass stream of pinned batches for ONE tower."""

    def __init__(self, dataset, batch_size: int, shuffle: bool = False, repeat: bool = True, rank: int = 0,
                 world: int = 1, workers: int = 4, prefetch: Optional[int] = None, seed: int = 0,
                 drop_last: bool = True, pin: Optional[bool] = None, device_augment: bool = False,
                 augment_fn: Optional[Callable] = None, device=None):
        """``device_augment``: the dataset must offer ``raw(i, epoch) -> (image_u8, mask_u8, params)``; batches are
        yielded as DEVICE tensors ([B,H,W,2] fp32, [B,H,W,1] fp32) produced by one augmentation kernel launch per
        batch on the consumer's current stream (``augment_fn`` defaults to ``gpu_augment.augment_batch``)."""
        self.ds, self.batch, self.shuffle, self.repeat = dataset, batch_size, shuffle, repeat
        self.device_augment, self.augment_fn, self.device = bool(device_augment), augment_fn, device
        if self.device_augment and not hasattr(dataset, "raw"):
            raise TypeError("device_augment=True needs a dataset with raw(i, epoch) -> (image_u8, mask_u8, params)")
        self.rank, self.world, self.workers, self.seed, self.drop_last = rank, world, max(workers, 1), seed, drop_last
        self.prefetch = prefetch or 2 * world
        self.pin = torch.cuda.is_available() if pin is None else pin

    def _index_stream(self) -> Iterator[Tuple[int, int]]:
        """Global sample stream (index, epoch) with a 10 x batch reservoir shuffle that crosses epochs (A16)."""
        rng = np.random.RandomState(self.seed)
        n = len(self.ds)
        buf: List[Tuple[int, int]] = []
        cap = self.batch * 10
        epoch = 0
        while True:
            for i in range(n):
                if not self.shuffle:
                    yield i, epoch
                    continue
                buf.append((i, epoch))
                if len(buf) >= cap:
                    j = rng.randint(len(buf))
                    buf[j], buf[-1] = buf[-1], buf[j]
                    yield buf.pop()
            epoch += 1
            if not self.repeat:
                break
        while buf:
            j = rng.randint(len(buf))
            buf[j], buf[-1] = buf[-1], buf[j]
            yield buf.pop()

    def _batches_for_this_tower(self) -> Iterator[List[Tuple[int, int]]]:
        """Tower r's batch of every COMPLETE round of ``world`` batches.  A single-pass stream whose batch count is
        not a multiple of ``world`` drops the ragged last round on every tower: otherwise the low ranks would run one
        more synchronous step than the others and wait for peers that never launch it."""
        cur, b, mine = [], 0, None
        for item in self._index_stream():
            cur.append(item)
            if len(cur) == self.batch:
                if b % self.world == self.rank:
                    mine = cur
                cur, b = [], b + 1
                if b % self.world == 0:
                    yield mine
                    mine = None
        if cur and not self.drop_last and self.world == 1:
            yield cur

    def _augment_on_device(self, xs: torch.Tensor, ys: torch.Tensor, params):
        """uint8 batch -> device (async copy from pinned memory) -> one augmentation kernel on the current stream."""
        fn = self.augment_fn
        if fn is None:
            from .gpu_augment import augment_batch as fn
        dev = self.device
        if dev is None:
            dev = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
        compat = getattr(self.ds, "reference_compat", True)
        return fn(xs.to(dev, non_blocking=True), ys.to(dev, non_blocking=True), params, reference_compat=compat)

    def __iter__(self):
        q: "queue.Queue" = queue.Queue(maxsize=self.prefetch)
        stop = threading.Event()

        def produce():
            from concurrent.futures import ThreadPoolExecutor

            try:
                with ThreadPoolExecutor(self.workers) as pool:
                    for idxs in self._batches_for_this_tower():
                        if stop.is_set():
                            return
                        if self.device_augment:  # decode only; the warp runs on the GPU, once per batch
                            raws = list(pool.map(lambda ie: self.ds.raw(ie[0], ie[1]), idxs))
                            xs = torch.from_numpy(np.stack([r[0] for r in raws]))
                            ys = torch.from_numpy(np.stack([r[1] for r in raws]))
                            params = [r[2] for r in raws]
                        else:
                            samples = list(pool.map(lambda ie: self.ds.__getitem__(ie[0], ie[1]), idxs))
                            xs, ys = _collate(samples)
                            params = None
                        if self.pin:
                            xs, ys = xs.pin_memory(), ys.pin_memory()
                        while not stop.is_set():
                            try:
                                q.put((xs, ys, params), timeout=0.1)
                                break
                            except queue.Full:
                                continue
                q.put(None)
            except Exception as e:  # noqa: BLE001
                q.put(e)

        t = threading.Thread(target=produce, daemon=True)
        t.start()
        try:
            while True:
                item = q.get()
                if item is None:
                    return
                if isinstance(item, Exception):
                    raise item
                xs, ys, params = item
                if params is not None:
                    yield self._augment_on_device(xs, ys, params)
                else:
                    yield xs, ys
        finally:
            stop.set()
