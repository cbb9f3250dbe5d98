"""Reference-compatible ``utils`` module (reference: utils.py)."""
from tensorflowdistributedlearning_b200.utils.misc import (get_available_gpus, metric_comparison,  # noqa: F401
                                                           metric_comparisson)
