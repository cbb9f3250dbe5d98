"""Training / evaluation hooks (replaces tf.train.*Hook used by the reference's Estimator run).

* ``CheckpointSaverHook`` -- save at step 0-equivalent (first call), every N steps and at the end
  (reference: RunConfig(save_checkpoints_steps=500), model.py:118; log Test.ipynb:203,207)
* ``StepCounterHook``     -- ``global_step/sec`` + loss every 100 steps (Test.ipynb:209-213), device-timed,
  max over towers
* ``SummarySaverHook``    -- scalars (and up to one image per tag: ``{mode}_image/_label/_prob/_prediction``,
  model.py:406-439, handed over in ``values["images"]``) every N steps into ``{dir}`` as TensorBoard
  events plus a JSONL mirror (reference: SummarySaverHook(save_steps=20 / 1), model.py:470-481)
* ``NanLossHook``         -- abort on a non-finite loss (failure detection)
Only tower 0 writes files (A27); every tower runs the hooks so collective calls stay aligned.
"""
from __future__ import annotations

import json
import math
import os
import time
from typing import Dict, Optional

import torch


class Hook:
    def begin(self, estimator) -> None: ...
    def after_step(self, estimator, step: int, values: Dict) -> None: ...
    def end(self, estimator, step: int) -> None: ...


class CheckpointSaverHook(Hook):
    def __init__(self, model_dir: str, save_steps: int = 500, keep: int = 5, listeners=()):
        self.model_dir, self.save_steps, self.keep = model_dir, save_steps, keep
        self.listeners = list(listeners)
        self._last = None

    def _save(self, estimator, step: int) -> None:
        if self._last == step:
            return
        self._last = step
        estimator.prepare_save()  # every tower: gathers the sharded optimizer slots so tower 0's copy is complete
        if estimator.group.is_primary:
            path = estimator.save_checkpoint(step, keep=self.keep)
        else:
            path = None
        estimator.group.barrier()
        for fn in self.listeners:
            fn(step, path)

    def begin(self, estimator) -> None:
        self._save(estimator, estimator.global_step)  # "Saving checkpoints for 0" (Test.ipynb:207)

    def after_step(self, estimator, step, values) -> None:
        if self.save_steps and step % self.save_steps == 0:
            self._save(estimator, step)

    def end(self, estimator, step) -> None:
        self._save(estimator, step)


class StepCounterHook(Hook):
    def __init__(self, every_n_steps: int = 100, log=print):
        self.every, self.log = every_n_steps, log
        self._t0 = None
        self._s0 = 0

    def begin(self, estimator) -> None:
        self._t0, self._s0 = time.perf_counter(), estimator.global_step

    def after_step(self, estimator, step, values) -> None:
        if self.every and step % self.every == 0:
            if estimator.group.device.type == "cuda":
                torch.cuda.synchronize()
            dt = estimator.group.max_over_ranks(time.perf_counter() - self._t0)
            rate = (step - self._s0) / max(dt, 1e-9)
            loss = values.get("loss")
            if estimator.group.is_primary:
                self.log(f"INFO:global_step/sec: {rate:.5g}  images/sec: {rate * estimator.global_batch:.1f}")
                self.log(f"INFO:loss = {loss:.7g}, step = {step} ({dt:.3f} sec)")
            self._t0, self._s0 = time.perf_counter(), step


class SummaryWriter:
    """TensorBoard event file (when tensorboard is importable) + ``summaries.jsonl`` in the same directory."""

    def __init__(self, logdir: str):
        os.makedirs(logdir, exist_ok=True)
        self.logdir = logdir
        self._jsonl = open(os.path.join(logdir, "summaries.jsonl"), "a")
        try:
            from torch.utils.tensorboard import SummaryWriter as TBWriter

            self._tb = TBWriter(logdir)
        except Exception:  # noqa: BLE001
            self._tb = None

    def scalars(self, step: int, values: Dict[str, float]) -> None:
        rec = {"step": int(step), **{k: float(v) for k, v in values.items()}}
        self._jsonl.write(json.dumps(rec) + "\n")
        self._jsonl.flush()
        if self._tb is not None:
            for k, v in values.items():
                self._tb.add_scalar(k, float(v), step)

    def image(self, tag: str, img_hw: torch.Tensor, step: int) -> None:
        """img_hw: [H, W] float tensor; written as an 8-bit PNG by the event writer (CPU encode, no graphics API).
        The JSONL mirror records that the tag was written (tag, step, shape)."""
        self._jsonl.write(json.dumps({"step": int(step), "image": tag, "shape": list(img_hw.shape)}) + "\n")
        self._jsonl.flush()
        if self._tb is None:
            return
        x = img_hw.detach().float().cpu()
        lo, hi = float(x.min()), float(x.max())
        x = (x - lo) / (hi - lo) if hi > lo else torch.zeros_like(x)
        self._tb.add_image(tag, x.unsqueeze(0), step)

    def close(self) -> None:
        self._jsonl.close()
        if self._tb is not None:
            self._tb.close()


class SummarySaverHook(Hook):
    def __init__(self, output_dir: str, save_steps: int = 20):
        self.output_dir, self.save_steps = output_dir, save_steps
        self._writer: Optional[SummaryWriter] = None

    def begin(self, estimator) -> None:
        if estimator.group.is_primary:
            self._writer = SummaryWriter(self.output_dir)

    def after_step(self, estimator, step, values) -> None:
        if self._writer is None or not self.save_steps or step % self.save_steps:
            return
        scalars = {k: v for k, v in values.items() if isinstance(v, (int, float))}
        self._writer.scalars(step, scalars)
        for tag, img in values.get("images", {}).items():
            self._writer.image(tag, img, step)

    def end(self, estimator, step) -> None:
        if self._writer is not None:
            self._writer.close()
            self._writer = None


class NanLossHook(Hook):
    def after_step(self, estimator, step, values) -> None:
        loss = values.get("loss")
        if loss is not None and not math.isfinite(loss):
            raise FloatingPointError(f"loss is {loss} at step {step} on tower {estimator.group.rank}")
