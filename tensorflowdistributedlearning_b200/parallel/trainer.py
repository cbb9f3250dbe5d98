"""One tower's training step: device-side input tail -> forward -> loss -> backward -> fused comm+optimizer,
captured once into a CUDA graph and replayed.

This is the hot loop that ``session.run(train_op)`` is for the reference, re-designed:
the loss stays on the GPU (the reference pins it to the CPU, model.py:391), the learning-rate schedule
and the Adam bias correction are device-resident, the whole step is one graph launch, the next batch's
host->device copy runs on a copy stream under the current step, and the only per-step host
synchronisation is the read of the scalar loss.
"""
from __future__ import annotations

import os

from typing import Callable, Optional, Tuple

import torch

from ..ops import functional as TF
from .optimizer import DistributedOptimizer
from .tower import TowerGroup


def classification_loss(logits: torch.Tensor, labels: torch.Tensor):
    loss, top1, top5 = TF.softmax_cross_entropy(logits.float(), labels)
    return loss, {"top1": top1, "top5": top5}


def segmentation_loss(logits: torch.Tensor, labels: torch.Tensor):
    loss = TF.lovasz_loss(labels, logits)
    iou, acc = TF.seg_metrics(logits, labels, 0.5, True)
    return loss, {"iou": iou, "acc": acc}


class TowerTrainer:
    def __init__(self, model: torch.nn.Module, group: TowerGroup, *, batch: int, input_hwc: Tuple[int, int, int],
                 label_shape: Tuple[int, ...] = (), label_dtype=torch.int64,
                 loss_fn: Callable = classification_loss, compute_dtype=torch.bfloat16, use_graph: bool = True,
                 mean=None, std=None, input_is_u8: bool = True, keep_output_sample: bool = False, **opt_kw):
        self.group = group
        dev = group.device
        self.device = dev
        self.model = model.to(dev)
        self.model.train()
        self.loss_fn = loss_fn
        self.compute_dtype = compute_dtype if dev.type == "cuda" else torch.float32
        self.opt = DistributedOptimizer(self.model, group, **opt_kw)
        h, w, c = input_hwc
        self.input_is_u8 = input_is_u8
        in_dtype = torch.uint8 if input_is_u8 else torch.float32
        self.images = torch.zeros((batch, h, w, c), dtype=in_dtype, device=dev)
        self.labels = torch.zeros((batch, *label_shape), dtype=label_dtype, device=dev)
        self.mean = torch.tensor(mean if mean is not None else [0.0] * c, dtype=torch.float32, device=dev)
        self.inv_std = 1.0 / torch.tensor(std if std is not None else [1.0] * c, dtype=torch.float32, device=dev)
        # `loss` is what gets logged: the MEAN over towers (EstimatorSpec(loss=...) under MirroredStrategy, model.py:491-500,
        # SURVEY X5); `local_loss` is this tower's own.  `metric_sums[k]` is the tower mean of the per-tower SUM of
        # metric k over the batch -- times the world size it is the global sum `tf.metrics.mean` accumulates (X4).
        # All of them ride behind the moving statistics in the packed one-shot reduction of the step.
        self.loss = torch.zeros((), dtype=torch.float32, device=dev)
        self.local_loss = torch.zeros((), dtype=torch.float32, device=dev)
        self.metrics = {}
        self.metric_sums = {}
        self.keep_output_sample = keep_output_sample
        self.output_sample: Optional[torch.Tensor] = None  # model output of sample 0 (image summaries, model.py:406-439)
        self.use_graph = bool(use_graph and dev.type == "cuda")
        self._nvtx = os.environ.get("TDL_NVTX", "0") == "1"
        if dev.type == "cuda":
            from ..ops import functional as TF

            # one launch per step re-lays every convolution weight (instead of one per layer): the trainer owns the
            # step boundaries this needs (TDL_BATCHED_WPREP=0 restores the per-layer launches)
            self._wprep = TF.enable_batched_weight_prep(os.environ.get("TDL_BATCHED_WPREP", "1") != "0")
        self._graph: Optional[torch.cuda.CUDAGraph] = None
        self._copy_stream = torch.cuda.Stream() if dev.type == "cuda" else None
        self._copied = torch.cuda.Event() if dev.type == "cuda" else None
        self._stage_free = torch.cuda.Event() if dev.type == "cuda" else None
        self._stage_images = None
        self._stage_labels = None
        self._staged = False
        self.launches_per_step = 0

    # ---- the step body (eager; also what gets captured) ----
    def _prepare_input(self) -> torch.Tensor:
        c_pad = getattr(self.model, "input_channel_pad", None) if self.device.type == "cuda" else None
        if self.device.type == "cuda" and self.input_is_u8 and self.compute_dtype == torch.bfloat16:
            from ..ops import native_ops as O

            return O.u8_to_bf16_normalize(self.images, self.mean, self.inv_std, c_pad=c_pad)
        x = self.images.float()
        if self.input_is_u8:
            x = x / 255.0
        x = ((x - self.mean) * self.inv_std).to(self.compute_dtype)
        if c_pad and c_pad > x.shape[-1]:
            x = torch.nn.functional.pad(x, (0, c_pad - x.shape[-1]))
        return x

    def _step_body(self) -> None:
        # NVTX ranges per phase (TDL_NVTX=1): eager steps only -- inside a captured graph the ranges would mark the
        # capture, not the replays (`scripts/profile_step.py` gives the per-kernel table)
        nvtx = self._nvtx and self.device.type == "cuda" and not torch.cuda.is_current_stream_capturing()
        mark = torch.cuda.nvtx.range_push if nvtx else (lambda *_: None)
        done = torch.cuda.nvtx.range_pop if nvtx else (lambda: None)
        if self.device.type == "cuda":
            from ..ops import functional as TF

            TF.begin_step()
        mark("tdl/input")
        x = self._prepare_input()
        done()
        self.opt.begin_step()
        mark("tdl/forward")
        out = self.model(x)
        loss, metrics = self.loss_fn(out, self.labels)
        done()
        mark("tdl/backward+bucket_allreduce")
        loss.backward()
        done()
        scalars = self.opt.flat.scalars
        keys = sorted(metrics)[:scalars.numel() - 1]
        self.local_loss.copy_(loss.detach())
        scalars[:1 + len(keys)].copy_(torch.stack([loss.detach().float()] +
                                                  [metrics[k].detach().float().sum() for k in keys]))
        mark("tdl/optimizer_tail")
        self.opt.step()  # gradients + optimizer, then ONE packed mean over towers: moving statistics, loss, metric sums
        done()
        self.loss.copy_(scalars[0])
        for i, k in enumerate(keys):
            if k not in self.metric_sums:
                self.metric_sums[k] = torch.zeros((), dtype=torch.float32, device=self.device)
            self.metric_sums[k].copy_(scalars[1 + i])
        for k, v in metrics.items():
            if k not in self.metrics:
                self.metrics[k] = torch.zeros_like(v)
            self.metrics[k].copy_(v)
        if self.keep_output_sample:
            if self.output_sample is None:
                self.output_sample = torch.zeros_like(out[0], dtype=torch.float32)
            self.output_sample.copy_(out[0].detach())
        if self.device.type == "cuda":
            TF.invalidate_prepared_weights()  # the optimizer has moved the weights

    def warmup_and_capture(self, warmup: int = 3) -> None:
        """Eager warm-up steps (they train: they are real steps), then capture one step into a CUDA graph."""
        from .. import _native as N

        if self.device.type != "cuda":
            return
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for i in range(max(warmup, 1)):
                before = N.kernel_launches
                self._step_body()
                self.launches_per_step = N.kernel_launches - before
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        if self.use_graph:
            self.group.barrier()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._step_body()
            self._graph = g
            torch.cuda.synchronize()
            self.group.barrier()

    def run_step(self) -> None:
        """One training step on whatever is in the device input buffers (no host traffic)."""
        if self._graph is not None:
            self._graph.replay()
            from ..ops import functional as TF

            TF.invalidate_prepared_weights()  # replays run no Python: say so for any eager forward that follows
        else:
            self._step_body()

    # ---- public end-to-end call ----
    def stage_batch(self, images_host: torch.Tensor, labels_host: torch.Tensor) -> None:
        """Asynchronously copy a (pinned) host batch into the device *staging* buffers on the copy stream."""
        if self.device.type != "cuda":
            self.images.copy_(images_host)
            self.labels.copy_(labels_host)
            self._staged = True
            return
        if self._stage_images is None:
            self._stage_images = torch.empty_like(self.images)
            self._stage_labels = torch.empty_like(self.labels)
        if images_host.is_cuda:
            # a batch the input pipeline already produced on the device (GPU augmentation): it was written on the
            # current stream, so the move into the staging buffers simply follows it there
            cur = torch.cuda.current_stream()
            self._stage_images.copy_(images_host, non_blocking=True)
            self._stage_labels.copy_(labels_host, non_blocking=True)
            self._copied.record(cur)
            self._staged = True
            return
        self._copy_stream.wait_event(self._stage_free)  # the previous staging contents have been consumed
        with torch.cuda.stream(self._copy_stream):
            self._stage_images.copy_(images_host, non_blocking=True)
            self._stage_labels.copy_(labels_host, non_blocking=True)
            self._copied.record(self._copy_stream)
        self._staged = True

    def train_step(self, images_host: torch.Tensor, labels_host: torch.Tensor, prefetch=None) -> float:
        """Host batch in, scalar loss out.

        The batch is copied host->device (pinned memory, copy stream), moved into the graph's input buffers,
        one training step runs, and the scalar loss is read back.  ``prefetch=(images, labels)`` starts the
        NEXT batch's host->device copy before this call blocks on the loss, so the copy hides under compute;
        the next call must then be given that same batch.
        """
        if not self._staged:
            self.stage_batch(images_host, labels_host)
        self._staged = False
        if self.device.type == "cuda":
            cur = torch.cuda.current_stream()
            cur.wait_event(self._copied)
            self.images.copy_(self._stage_images, non_blocking=True)
            self.labels.copy_(self._stage_labels, non_blocking=True)
            self._stage_free.record(cur)
        self.run_step()
        if prefetch is not None:
            self.stage_batch(*prefetch)
        try:
            return float(self.loss.item())
        except RuntimeError as e:
            who = self.opt.peer_timeout()
            if who is not None:
                raise RuntimeError(f"tower {who['rank']} gave up waiting for tower {who['peer']} in the gradient "
                                   f"exchange (CTA {who['cta']}, epoch {who['epoch']}): that tower died, hung, or is "
                                   f"more than the collective timeout behind") from e
            raise

    def close(self) -> None:
        if self.device.type == "cuda":
            from ..ops import functional as TF

            TF.enable_batched_weight_prep(False)  # drops the persistent copies and the table
        self._graph = None
        self.opt.close()
