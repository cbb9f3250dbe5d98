"""Small utilities (reference: utils.py)."""
from __future__ import annotations

from typing import Dict, List

import torch


def get_available_gpus() -> List[str]:
    """Names of the local GPU devices (reference: utils.py:6-8)."""
    return [f"/device:GPU:{i}" for i in range(torch.cuda.device_count())] if torch.cuda.is_available() else []


def metric_comparison(best_eval_result: Dict, current_eval_result: Dict, key: str,
                      greater_is_better: bool = True, reference_compat: bool = False) -> bool:
    """True when ``current`` is better than ``best`` on ``key`` -- what BestExporter.compare_fn must return.

    The reference's ``metric_comparisson`` (utils.py:11-28) returns ``best > current`` for
    ``greater_is_better``, i.e. it exports when the model got worse (defect 4);
    ``reference_compat=True`` reproduces that.
    """
    if not best_eval_result or key not in best_eval_result:
        raise ValueError(f"best_eval_result cannot be empty or no {key} metric is found in it.")
    if not current_eval_result or key not in current_eval_result:
        raise ValueError(f"current_eval_result cannot be empty or no {key} metric is found in it.")
    b, c = best_eval_result[key], current_eval_result[key]
    if reference_compat:
        return b > c if greater_is_better else b < c
    return c > b if greater_is_better else c < b


metric_comparisson = metric_comparison  # the reference's spelling
