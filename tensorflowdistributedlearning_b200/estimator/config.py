"""Configuration dataclasses (replaces tf.estimator.RunConfig / TrainSpec / EvalSpec and the reference's
module-level constants + **kwargs, model.py:13-24,29-38,63-121)."""
from __future__ import annotations

from dataclasses import asdict, dataclass, field
from typing import Any, Callable, Dict, Optional, Tuple


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
