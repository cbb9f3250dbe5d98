"""Configuration dataclasses (replaces tf.estimator.RunConfig / TrainSpec / EvalSpec and the reference's
module-level constants + **kwargs, model.py:13-24,29-38,63-121)."""
from __future__ import annotations

import ast
from dataclasses import asdict, dataclass, field, fields
from typing import Any, Callable, Dict, Iterable, Optional, Tuple


@dataclass
class ModelConfig:
    name: str = "refseg"
    weight_decay: float = 0.001          # model.py:14
    batch_norm_decay: float = 0.99       # model.py:16
    batch_norm_epsilon: float = 0.001    # model.py:17
    batch_norm_scale: bool = True        # model.py:18
    output_stride: int = 8               # model.py:20
    input_shape: Tuple[int, int] = (101, 101)  # model.py:22
    base_depth: int = 256                # model.py:24
    n_blocks: Tuple[int, ...] = (3, 4, 6)
    block_type: str = "bottleneck"
    in_channels: int = 2
    num_classes: Optional[int] = None

    def build(self, model_name: str = "model"):
        """The network this configuration names: ``refseg`` = the reference's segmentation network under the variable
        scope ``model_name`` (model.py:353-370); anything else goes through ``models.build``."""
        from .. import models

        if self.name.lower().replace("_", "") in ("refseg", "resnetmodel", "seg"):
            return models.ResNetSegModel(
                model_name=model_name, in_channels=self.in_channels, weight_decay=self.weight_decay,
                batch_norm_decay=self.batch_norm_decay, batch_norm_epsilon=self.batch_norm_epsilon,
                batch_norm_scale=self.batch_norm_scale, output_stride=self.output_stride, base_depth=self.base_depth,
                input_shape=tuple(self.input_shape), n_blocks=tuple(self.n_blocks), block_type=self.block_type)
        kw = {} if self.num_classes is None else {"num_classes": self.num_classes}
        return models.build(self.name, **kw)


@dataclass
class OptimConfig:
    kind: str = "adam"                   # model.py:462
    lr: float = 0.001                    # model.py:32
    decay_steps: float = 10000.0         # model.py:457-458
    decay_rate: float = 0.5
    momentum: float = 0.9
    weight_decay: float = 0.0            # the reference's L2 regulariser never reaches the loss
    beta1: float = 0.9
    beta2: float = 0.999
    eps: float = 1e-8
    comm: Optional[str] = None           # p2p | nccl | gloo | None (auto)
    bucket_cap_mb: float = 32.0
    overlap: bool = True


@dataclass
class RunConfig:
    model_dir: str = "model"
    save_checkpoints_steps: int = 500    # model.py:118
    keep_checkpoint_max: int = 5         # TF default, Test.ipynb:192
    save_summary_steps: int = 20         # SummarySaverHook(save_steps=20), model.py:470-473
    log_step_count_steps: int = 100      # Test.ipynb:192
    seed: int = 42
    use_cuda_graph: bool = True
    nan_check: bool = True

    def to_dict(self) -> Dict[str, Any]:
        return asdict(self)


@dataclass
class TrainSpec:
    input_fn: Callable[[], Any]          # () -> iterator of (images, labels) host batches for THIS tower
    max_steps: int = 100


@dataclass
class EvalSpec:
    input_fn: Callable[[], Any]
    steps: Optional[int] = None
    throttle_secs: int = 300             # model.py:214
    start_delay_secs: int = 0
    exporters: Tuple = field(default_factory=tuple)
    name: str = "eval"
    sharded: bool = False                # True: input_fn yields this tower's share, all towers evaluate (SUM on read)


def apply_overrides(sections: Dict[str, Any], items: Iterable[str]) -> Dict[str, Any]:
    """Command-line overrides of dataclass fields: every item is ``section.field=value`` (``model.base_depth=128``,
    ``optim.lr=0.01``, ``run.save_checkpoints_steps=100``); values are Python literals, bare words stay strings.
    Unknown sections / fields are errors ("a CLI that overrides dataclass fields")."""
    for item in items or ():
        if "=" not in item or "." not in item.split("=", 1)[0]:
            raise ValueError(f"override '{item}': expected section.field=value")
        key, raw = item.split("=", 1)
        section, name = key.split(".", 1)
        if section not in sections:
            raise ValueError(f"override '{item}': unknown section '{section}' (have {sorted(sections)})")
        cfg = sections[section]
        if name not in {f.name for f in fields(cfg)}:
            raise ValueError(f"override '{item}': {type(cfg).__name__} has no field '{name}'")
        try:
            value = ast.literal_eval(raw)
        except (ValueError, SyntaxError):
            value = raw
        cur = getattr(cfg, name)
        if isinstance(cur, tuple) and isinstance(value, list):
            value = tuple(value)
        if isinstance(cur, float) and isinstance(value, int) and not isinstance(value, bool):
            value = float(value)
        setattr(cfg, name, value)
    return sections
