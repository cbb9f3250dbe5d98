"""Reference-compatible ``core.metric`` (reference: core/metric.py:6-71): the reference's call signatures.

``mIOU(y_true, y_pred, metrics_collections=None, updates_collections=None, name="iou") -> (value, update_op)`` and
``mean_accuracy(...)`` keep their streaming accumulators per ``name`` (TF local variables);
``reset_local_variables()`` re-initialises them."""
from tensorflowdistributedlearning_b200.compat import mIOU, mean_accuracy, reset_local_variables  # noqa: F401
from tensorflowdistributedlearning_b200.ops.reference import IOU_THRESHOLDS  # noqa: F401
