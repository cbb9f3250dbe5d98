"""Estimator: train / evaluate / predict / train_and_evaluate over towers (replaces tf.estimator usage in
model.py:117-121,164-223,245-255 and the `_model_fn` orchestration of model.py:326-505).

Behavioural contract kept from the reference:
  * training resumes from the newest checkpoint in ``model_dir``; checkpoints every ``save_checkpoints_steps``
  * train metrics are streaming means since the start of training; eval metrics restart per evaluate call
  * ``train_and_evaluate``: train continuously; after a checkpoint, if >= ``throttle_secs`` since the last
    evaluation, evaluate that state and hand the result to the exporters; final evaluate + export at max_steps
  * prediction runs on tower 0's device; evaluation is sharded over the towers by default (every tower evaluates
    batches r, r+N, ... of the same stream and the metric accumulators are SUM-reduced on read -- same numbers as the
    reference's single-device evaluation, and no tower idles behind a timed barrier while tower 0 evaluates)
  * summaries / checkpoints are written by tower 0; the logged loss is the tower mean (it rides, with the metric
    sums, in the packed statistics reduction of the step)
"""
from __future__ import annotations

import json
import os
import time
from typing import Callable, Dict, Iterable, Iterator, List, Optional

import torch

from ..data import preprocessing as PP
from ..ops import functional as TF
from ..parallel.tower import TowerGroup
from ..parallel.trainer import TowerTrainer, classification_loss, segmentation_loss
from . import checkpoint as ckpt
from .config import EvalSpec, OptimConfig, RunConfig, TrainSpec
from .hooks import CheckpointSaverHook, Hook, NanLossHook, StepCounterHook, SummarySaverHook, SummaryWriter


class StreamingMean:
    """tf.metrics.mean: total += sum(v), count += numel(v); value = total / count (tower-local, SUM on read)."""

    def __init__(self, device):
        self.total = torch.zeros((), dtype=torch.float64, device=device)
        self.count = torch.zeros((), dtype=torch.float64, device=device)

    def update(self, v: torch.Tensor) -> None:
        self.total += v.detach().double().sum()
        self.count += v.numel()

    def update_global(self, total: torch.Tensor, count: float) -> None:
        """Add an already cross-tower-reduced (sum, count): the accumulators then hold the global totals on every
        tower and ``value()`` needs no collective."""
        self.total += total.detach().double()
        self.count += count

    def value(self, group: Optional[TowerGroup] = None) -> float:
        t = torch.stack([self.total, self.count])
        if group is not None and group.distributed:
            t = t.clone()
            group.all_reduce_sum_(t)
        return float(t[0] / t[1]) if float(t[1]) > 0 else 0.0

    def reset(self) -> None:
        self.total.zero_()
        self.count.zero_()


TASKS = {
    "segmentation": dict(loss_fn=segmentation_loss, metric_keys={"iou": "metrics/mean_iou", "acc": "metrics/mean_acc"},
                         loss_key="loss/lovasz_loss"),
    "classification": dict(loss_fn=classification_loss, metric_keys={"top1": "metrics/top1", "top5": "metrics/top5"},
                           loss_key="loss/softmax_cross_entropy"),
}


class Estimator:
    def __init__(self, model_builder: Callable[[], torch.nn.Module], model_dir: str, config: Optional[RunConfig] = None,
                 optim: Optional[OptimConfig] = None, params: Optional[dict] = None, group: Optional[TowerGroup] = None,
                 task: str = "segmentation", input_is_u8: bool = False, mean=None, std=None):
        self.model_builder = model_builder
        self.model_dir = model_dir
        self.config = config or RunConfig(model_dir=model_dir)
        self.optim = optim or OptimConfig()
        self.params = dict(params or {})
        self.group = group or TowerGroup()
        self.task = TASKS[task]
        self.task_name = task
        self.input_is_u8, self.mean, self.std = input_is_u8, mean, std
        self.model: Optional[torch.nn.Module] = None
        self.trainer: Optional[TowerTrainer] = None
        self.global_batch = 0
        self.n_params: Optional[int] = None
        self._train_metrics: Dict[str, StreamingMean] = {}
        if self.group.is_primary:
            os.makedirs(model_dir, exist_ok=True)

    # ---- model / trainer construction ----
    def _ensure_model(self) -> torch.nn.Module:
        if self.model is None:
            torch.manual_seed(self.config.seed)
            self.model = self.model_builder().to(self.group.device)
            # reference counts ALL variables in scope, moving statistics included (model.py:444-445)
            from ..ops.layers import iter_tf_variables

            n = sum(t.numel() for _, t, _ in iter_tf_variables(self.model))
            self.n_params = n or (sum(p.numel() for p in self.model.parameters()) + sum(
                b.numel() for b in self.model.buffers() if b.is_floating_point()))
        return self.model

    def _ensure_trainer(self, images: torch.Tensor, labels: torch.Tensor) -> TowerTrainer:
        if self.trainer is None:
            model = self._ensure_model()
            o = self.optim
            self.trainer = TowerTrainer(
                model, self.group, batch=images.shape[0], input_hwc=tuple(images.shape[1:]),
                label_shape=tuple(labels.shape[1:]), label_dtype=labels.dtype, loss_fn=self.task["loss_fn"],
                use_graph=self.config.use_cuda_graph, input_is_u8=self.input_is_u8, mean=self.mean, std=self.std,
                keep_output_sample=self.task_name == "segmentation",
                kind=o.kind, lr=o.lr, momentum=o.momentum, weight_decay=o.weight_decay, beta1=o.beta1, beta2=o.beta2,
                eps=o.eps, decay_steps=o.decay_steps, decay_rate=o.decay_rate, comm=o.comm,
                bucket_cap_mb=o.bucket_cap_mb, overlap=o.overlap)
            self.global_batch = images.shape[0] * self.group.world
            latest = ckpt.latest_checkpoint(self.model_dir)
            if latest is not None:
                step = ckpt.load_checkpoint(latest, model, self.trainer.opt)
                if self.group.is_primary:
                    print(f"INFO:Restoring parameters from {latest} (step {step})")
            if self.group.device.type == "cuda":
                self.trainer.images.copy_(images)
                self.trainer.labels.copy_(labels)
                # warm-up steps are real training steps; keep the step counter honest by restoring afterwards
                snap = self._snapshot()
                self.trainer.warmup_and_capture(warmup=2)
                self._restore(snap)
        return self.trainer

    def _snapshot(self):
        t = self.trainer
        return (t.opt.flat.param_flat.clone(), t.opt.flat.stats_flat.clone(), t.opt.m1.clone(),
                None if t.opt.m2 is None else t.opt.m2.clone(), t.opt.global_step.clone())

    def _restore(self, snap) -> None:
        t = self.trainer
        t.opt.flat.param_flat.copy_(snap[0])
        t.opt.flat.stats_flat.copy_(snap[1])
        t.opt.m1.copy_(snap[2])
        if snap[3] is not None:
            t.opt.m2.copy_(snap[3])
        t.opt.global_step.copy_(snap[4])
        t.opt.flat.grad_flat.zero_()
        if self.group.device.type == "cuda":
            torch.cuda.synchronize()
        self.group.barrier()

    @property
    def global_step(self) -> int:
        if self.trainer is not None:
            return int(self.trainer.opt.global_step.item())
        latest = ckpt.latest_checkpoint(self.model_dir)
        if latest is None:
            return 0
        with open(latest + ".json") as f:
            return int(json.load(f)["global_step"])

    def prepare_save(self) -> None:
        """Collective part of a checkpoint: all towers call it before tower 0 writes (optimizer slots are sharded over
        the towers on the P2P two-shot path)."""
        if self.trainer is not None:
            self.trainer.opt.consolidate_state()

    def save_checkpoint(self, step: int, keep: int = 5) -> str:
        return ckpt.save_checkpoint(self.model_dir, self.model, self.trainer.opt if self.trainer else None, step, keep,
                                    extra={"params": {k: str(v) for k, v in self.params.items()}})

    # ---- train ----
    def train(self, input_fn: Callable[[], Iterable], max_steps: int, hooks: Optional[List[Hook]] = None,
              saving_listeners=()) -> "Estimator":
        it = iter(input_fn())
        try:
            images, labels = next(it)
        except StopIteration:
            return self
        trainer = self._ensure_trainer(images, labels)
        fold_dir = self.model_dir
        hooks = list(hooks or []) + [
            CheckpointSaverHook(self.model_dir, self.config.save_checkpoints_steps, self.config.keep_checkpoint_max,
                                saving_listeners),
            StepCounterHook(self.config.log_step_count_steps),
            SummarySaverHook(os.path.join(fold_dir, "train"), self.config.save_summary_steps),
        ]
        if self.config.nan_check:
            hooks.append(NanLossHook())
        for k in self.task["metric_keys"]:
            self._train_metrics.setdefault(k, StreamingMean(self.group.device))
        step = self.global_step
        if step >= max_steps:
            return self
        for h in hooks:
            h.begin(self)
        cur = (images, labels)
        while step < max_steps:
            try:
                nxt = next(it)
            except StopIteration:
                nxt = None
            loss = trainer.train_step(cur[0], cur[1], prefetch=nxt)
            step += 1
            world = self.group.world
            for k, m in self._train_metrics.items():
                if k in trainer.metric_sums:  # tower mean of the per-tower sums, reduced inside the step (X4)
                    m.update_global(trainer.metric_sums[k] * world, trainer.metrics[k].numel() * world)
                elif k in trainer.metrics:
                    m.update(trainer.metrics[k])
            values: Dict = {"loss": loss, self.task["loss_key"]: loss, "learning_rate": None}
            if self.config.save_summary_steps and step % self.config.save_summary_steps == 0:
                for k, name in self.task["metric_keys"].items():
                    reduced = k in trainer.metric_sums
                    values[name] = self._train_metrics[k].value(None if reduced else self.group)
                values["learning_rate"] = trainer.opt.current_lr()
                if self.group.is_primary and trainer.output_sample is not None:
                    # the reference's four train-mode image summaries, first sample only (model.py:406-439,470-473)
                    prob = torch.sigmoid(trainer.output_sample[:, :, 0])
                    values["images"] = {
                        "train_image": trainer.images[0, :, :, 0].float(), "train_label": trainer.labels[0, :, :, 0].float(),
                        "train_prob": prob, "train_prediction": (prob > self.params.get("threshold", 0.5)).float()}
            values = {k: v for k, v in values.items() if v is not None}
            for h in hooks:
                h.after_step(self, step, values)
            if nxt is None:
                break
            cur = nxt
        for h in hooks:
            h.end(self, step)
        return self

    # ---- evaluate ----
    @torch.no_grad()
    def evaluate(self, input_fn: Callable[[], Iterable], steps: Optional[int] = None, name: str = "eval",
                 checkpoint_path: Optional[str] = None, sharded: bool = False) -> Dict[str, float]:
        """``sharded=True`` (all towers must call): ``input_fn`` yields THIS tower's batches (r, r+N, ... of the global
        stream), ``steps`` counts global batches, and the accumulators are SUM-reduced over towers when read."""
        model = self._ensure_model()
        sharded = bool(sharded and self.group.distributed)
        if sharded and steps is not None:
            steps = len(range(self.group.rank, steps, self.group.world))
        if checkpoint_path is not None:
            ckpt.load_checkpoint(checkpoint_path, model, None)
        elif self.trainer is None:
            latest = ckpt.latest_checkpoint(self.model_dir)
            if latest is not None:
                ckpt.load_checkpoint(latest, model, None)
        was_training = model.training
        model.eval()
        dev = self.group.device
        metrics = {k: StreamingMean(dev) for k in self.task["metric_keys"]}
        loss_mean = StreamingMean(dev)
        writer = SummaryWriter(os.path.join(self.model_dir, name)) if self.group.is_primary else None
        n = 0
        last = None
        for images, labels in input_fn():
            if steps is not None and n >= steps:
                break
            x = self._to_device_input(images)
            y = labels.to(dev, non_blocking=True)
            logits = model(x)
            loss, m = self.task["loss_fn"](logits, y)
            loss_mean.update(loss.reshape(1))
            for k in metrics:
                metrics[k].update(m[k])
            last = (x, y, logits)
            n += 1
        step = self.global_step
        g = self.group if sharded else None
        result = {name_: metrics[k].value(g) for k, name_ in self.task["metric_keys"].items()}
        result[self.task["loss_key"]] = loss_mean.value(g)
        result["loss"] = result[self.task["loss_key"]]
        result["global_step"] = step
        if writer is not None:
            writer.scalars(step, {k: v for k, v in result.items() if k != "global_step"})
            if last is not None and self.task_name == "segmentation":
                x, y, logits = last
                prob = torch.sigmoid(logits.float())
                writer.image("eval_image", x[0, :, :, 0], step)
                writer.image("eval_label", y[0, :, :, 0], step)
                writer.image("eval_prob", prob[0, :, :, 0], step)
                writer.image("eval_prediction", (prob[0, :, :, 0] > self.params.get("threshold", 0.5)).float(), step)
            writer.close()
        if was_training:
            model.train()
        return result

    def _to_device_input(self, images: torch.Tensor) -> torch.Tensor:
        dev = self.group.device
        x = images.to(dev, non_blocking=True)
        if self.input_is_u8:
            c = x.shape[-1]
            mean = torch.tensor(self.mean or [0.0] * c, device=dev)
            std = torch.tensor(self.std or [1.0] * c, device=dev)
            x = (x.float() / 255.0 - mean) / std
        return x.to(torch.bfloat16 if dev.type == "cuda" else torch.float32)

    # ---- predict ----
    @torch.no_grad()
    def predict(self, input_fn: Callable[[], Iterable], checkpoint_path: Optional[str] = None) -> Iterator[Dict]:
        """Yields per-batch {'probabilities', 'mask'}; undoes ``params['transformation']`` (TTA) on the outputs
        (reference: model.py:379-387,493-494)."""
        model = self._ensure_model()
        path = checkpoint_path or (None if self.trainer is not None else ckpt.latest_checkpoint(self.model_dir))
        if path is not None:
            ckpt.load_checkpoint(path, model, None)
        model.eval()
        tfm = self.params.get("transformation", "none")
        thr = self.params.get("threshold", 0.5)
        for batch in input_fn():
            images = batch[0] if isinstance(batch, (tuple, list)) else batch
            logits = model(self._to_device_input(images)).float()
            if self.task_name == "segmentation":
                prob = torch.sigmoid(logits)
                prob = PP.undo_transformation(prob, tfm)
                yield {"probabilities": prob.cpu(), "mask": (prob > thr).float().cpu()}
            else:
                yield {"probabilities": torch.softmax(logits, -1).cpu(), "classes": logits.argmax(-1).cpu()}

    # ---- train_and_evaluate ----
    def train_and_evaluate(self, train_spec: TrainSpec, eval_spec: EvalSpec) -> Dict[str, float]:
        state = {"last_eval": time.time() - eval_spec.throttle_secs + eval_spec.start_delay_secs, "result": None,
                 "evaluated_step": -1}

        def run_eval(step: int) -> None:
            if getattr(eval_spec, "sharded", False) and self.group.distributed:
                # every tower evaluates its share; nobody waits behind a collective timeout for tower 0
                res = self.evaluate(eval_spec.input_fn, eval_spec.steps, eval_spec.name, sharded=True)
            elif self.group.is_primary:
                res = self.evaluate(eval_spec.input_fn, eval_spec.steps, eval_spec.name)
            else:
                res = None
            state["result"] = res
            if self.group.is_primary:
                for ex in eval_spec.exporters:
                    ex.export(self, res)
            state["evaluated_step"] = step
            state["last_eval"] = time.time()
            self.group.barrier()

        def listener(step: int, path) -> None:
            due = time.time() - state["last_eval"] >= eval_spec.throttle_secs
            flag = torch.tensor([1.0 if due else 0.0], device=self.group.device)
            self.group.broadcast_(flag, 0)  # tower 0's clock decides so every tower takes the same branch
            if float(flag.item()) > 0 and step > 0:
                run_eval(step)

        self.train(train_spec.input_fn, train_spec.max_steps, saving_listeners=(listener,))
        if state["evaluated_step"] != self.global_step:
            run_eval(self.global_step)  # final evaluation + export
        return state["result"] or {}


class BestExporter:
    """Keep the ``exports_to_keep`` best models by ``compare_fn(best_eval_result, current_eval_result)`` (True ==
    current is better; A26).  The best result persists across restarts in ``best_eval.json``."""

    def __init__(self, name: str = "best_exporter", compare_fn: Optional[Callable] = None, exports_to_keep: int = 5):
        from ..utils.misc import metric_comparison

        self.name = name
        self.compare_fn = compare_fn or (lambda best, cur: metric_comparison(best, cur, "loss", False))
        self.keep = exports_to_keep

    def export(self, estimator: Estimator, eval_result: Dict[str, float]) -> Optional[str]:
        base = os.path.join(estimator.model_dir, "export", self.name)
        os.makedirs(base, exist_ok=True)
        best_file = os.path.join(base, "best_eval.json")
        best = None
        if os.path.exists(best_file):
            with open(best_file) as f:
                best = json.load(f)
        if best is not None and not self.compare_fn(best, eval_result):
            return None
        stamp = str(int(time.time() * 1000))
        out = os.path.join(base, stamp)
        os.makedirs(out, exist_ok=True)
        ckpt.save_checkpoint(out, estimator.model, None, int(eval_result.get("global_step", 0)), keep=1,
                             extra={"eval_result": eval_result, "serving_input_key": "images"})
        with open(best_file, "w") as f:
            json.dump(eval_result, f)
        exports = sorted(d for d in os.listdir(base) if d.isdigit())
        for old in exports[:-self.keep] if len(exports) > self.keep else []:
            import shutil

            shutil.rmtree(os.path.join(base, old), ignore_errors=True)
        return out
