from .misc import get_available_gpus, metric_comparison, metric_comparisson  # noqa: F401
