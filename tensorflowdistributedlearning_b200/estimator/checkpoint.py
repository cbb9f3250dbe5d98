"""Checkpoint I/O keyed by the reference's TF variable names.

A checkpoint is ``model.ckpt-{step}.safetensors`` (tensors in TF layout: conv weights HWIO, depthwise
[kh,kw,C,1]) + ``model.ckpt-{step}.json`` (manifest: names, shapes, dtypes, global_step), and a TF-style
``checkpoint`` state file naming the latest.  Optimizer slots use TF's names: ``<var>/Adam``, ``<var>/Adam_1``
(or ``<var>/Momentum``), ``beta1_power``, ``beta2_power``, ``global_step``.  Tower 0 writes (X7); every tower reads.

TensorFlow's own V2 format is supported as well (``estimator/tf_bundle.py``): ``load_checkpoint`` accepts a prefix
whose files are ``<prefix>.index`` + ``<prefix>.data-00000-of-00001`` (e.g. a checkpoint written by the reference),
and ``save_checkpoint(..., tf_bundle=True)`` writes that pair next to the safetensors file.
"""
from __future__ import annotations

import json
import os
import re
from typing import Dict, List, Optional

import torch

from ..ops.layers import _unpad, from_tf_layout, iter_tf_variables, iter_tf_variables_ex, to_tf_layout

STATE_FILE = "checkpoint"


def _ckpt_prefix(model_dir: str, step: int) -> str:
    return os.path.join(model_dir, f"model.ckpt-{step}")


def list_checkpoints(model_dir: str) -> List[int]:
    steps = []
    if os.path.isdir(model_dir):
        for f in os.listdir(model_dir):
            m = re.fullmatch(r"model\.ckpt-(\d+)\.(?:safetensors|index)", f)
            if m:
                steps.append(int(m.group(1)))
    return sorted(set(steps))


def latest_checkpoint(model_dir: str) -> Optional[str]:
    state = os.path.join(model_dir, STATE_FILE)
    if os.path.exists(state):
        with open(state) as f:
            m = re.search(r'model_checkpoint_path: "([^"]+)"', f.read())
        if m:
            p = m.group(1)
            p = p if os.path.isabs(p) else os.path.join(model_dir, p)
            if os.path.exists(p + ".safetensors") or os.path.exists(p + ".index"):
                return p
    steps = list_checkpoints(model_dir)
    return _ckpt_prefix(model_dir, steps[-1]) if steps else None


def model_tf_state(model: torch.nn.Module) -> Dict[str, torch.Tensor]:
    return {name: to_tf_layout(t.detach(), kind).float().cpu().contiguous() for name, t, kind in iter_tf_variables(model)}


def save_checkpoint(model_dir: str, model: torch.nn.Module, optimizer=None, step: Optional[int] = None,
                    keep: int = 5, extra: Optional[dict] = None, tf_bundle: bool = False) -> str:
    from safetensors.torch import save_file

    os.makedirs(model_dir, exist_ok=True)
    tensors = model_tf_state(model)
    if optimizer is not None:
        sd = optimizer.state_dict()
        step = sd["global_step"] if step is None else step
        flat = optimizer.flat
        slot1, slot2 = ("Adam", "Adam_1") if sd["kind"] == "adam" else ("Momentum", None)
        by_id = {id(t): (name, kind, shape) for name, t, kind, shape in iter_tf_variables_ex(optimizer.flat.model)}
        for p in flat.params:
            if id(p) not in by_id:
                continue
            name, kind, shape = by_id[id(p)]
            o = flat.offsets[id(p)]
            tensors[f"{name}/{slot1}"] = to_tf_layout(_unpad(sd["m1"][o:o + p.numel()].view(p.shape), shape),
                                                      kind).float().cpu().contiguous()
            if slot2 and sd["m2"] is not None:
                tensors[f"{name}/{slot2}"] = to_tf_layout(_unpad(sd["m2"][o:o + p.numel()].view(p.shape), shape),
                                                          kind).float().cpu().contiguous()
        if sd["kind"] == "adam":
            # TF initialises beta*_power to beta and multiplies once per apply: after `step` updates it holds
            # beta ** (step + 1) (so the step-0 checkpoint stores beta, never 1.0 -- 1 - beta1_power must be > 0)
            tensors["beta1_power"] = torch.tensor(optimizer.beta1 ** (step + 1), dtype=torch.float32)
            tensors["beta2_power"] = torch.tensor(optimizer.beta2 ** (step + 1), dtype=torch.float32)
    step = int(step or 0)
    tensors["global_step"] = torch.tensor(step, dtype=torch.int64)
    prefix = _ckpt_prefix(model_dir, step)
    save_file(tensors, prefix + ".safetensors.tmp")
    os.replace(prefix + ".safetensors.tmp", prefix + ".safetensors")
    if tf_bundle:  # the same variables as a TensorFlow V2 checkpoint (readable by tf.train.Saver / the reference)
        from .tf_bundle import write_bundle

        write_bundle(prefix, {k: v.numpy() for k, v in tensors.items()})
    manifest = {"global_step": step, "variables": {k: {"shape": list(v.shape), "dtype": str(v.dtype)} for k, v in tensors.items()}}
    if extra:
        manifest["extra"] = extra
    with open(prefix + ".json", "w") as f:
        json.dump(manifest, f, indent=1)
    steps = list_checkpoints(model_dir)
    for old in steps[:-keep] if keep and len(steps) > keep else []:
        for ext in (".safetensors", ".json", ".index", ".data-00000-of-00001"):
            try:
                os.unlink(_ckpt_prefix(model_dir, old) + ext)
            except OSError:
                pass
    steps = list_checkpoints(model_dir)
    with open(os.path.join(model_dir, STATE_FILE), "w") as f:
        f.write(f'model_checkpoint_path: "model.ckpt-{step}"\n')
        for s in steps:
            f.write(f'all_model_checkpoint_paths: "model.ckpt-{s}"\n')
    return prefix


def load_checkpoint(prefix: str, model: torch.nn.Module, optimizer=None, strict: bool = True) -> int:
    """Restore weights, moving statistics, optimizer slots and global_step.  Returns the step."""
    if os.path.exists(prefix + ".safetensors"):
        from safetensors.torch import load_file

        tensors = load_file(prefix + ".safetensors")
    else:  # a TensorFlow V2 checkpoint (<prefix>.index + <prefix>.data-*), e.g. one trained with the reference
        from .tf_bundle import read_bundle

        tensors = {k: torch.from_numpy(v) for k, v in read_bundle(prefix).items()}
    missing = []
    with torch.no_grad():
        for name, t, kind in iter_tf_variables(model):
            if name not in tensors:
                missing.append(name)
                continue
            t.copy_(from_tf_layout(tensors[name], kind).to(t.device, t.dtype))
    if strict and missing:
        raise KeyError(f"checkpoint {prefix} lacks {len(missing)} variables, e.g. {missing[:3]}")
    step = int(tensors["global_step"].item()) if "global_step" in tensors else 0
    if optimizer is not None:
        flat = optimizer.flat
        slot1, slot2 = ("Adam", "Adam_1") if optimizer.kind == "adam" else ("Momentum", None)
        by_id = {id(t): (name, kind, shape) for name, t, kind, shape in iter_tf_variables_ex(flat.model)}
        with torch.no_grad():
            for p in flat.params:
                if id(p) not in by_id:
                    continue
                name, kind, shape = by_id[id(p)]
                o = flat.offsets[id(p)]
                if f"{name}/{slot1}" in tensors:
                    _unpad(optimizer.m1[o:o + p.numel()].view(p.shape), shape).copy_(
                        from_tf_layout(tensors[f"{name}/{slot1}"], kind))
                if slot2 and optimizer.m2 is not None and f"{name}/{slot2}" in tensors:
                    _unpad(optimizer.m2[o:o + p.numel()].view(p.shape), shape).copy_(
                        from_tf_layout(tensors[f"{name}/{slot2}"], kind))
            optimizer.global_step.fill_(step)
    return step
