"""Reference-compatible ``core.metric`` (reference: core/metric.py): streaming mIOU and pixel accuracy."""
import torch

from tensorflowdistributedlearning_b200.estimator.estimator import StreamingMean
from tensorflowdistributedlearning_b200.ops.functional import seg_metrics
from tensorflowdistributedlearning_b200.ops.reference import IOU_THRESHOLDS  # noqa: F401


class _Streaming:
    def __init__(self, which, reference_compat=True):
        self.which, self.compat, self.mean = which, reference_compat, None

    def update(self, y_true, y_pred_logits):
        iou, acc = seg_metrics(y_pred_logits, y_true, 0.5, self.compat)
        v = iou if self.which == "iou" else acc
        if self.mean is None:
            self.mean = StreamingMean(v.device)
        self.mean.update(v)
        return self.mean.value()

    def value(self):
        return 0.0 if self.mean is None else self.mean.value()


def mIOU(reference_compat=True):
    """-> object with ``update(y_true, logits)`` (the TF ``update_op``) and ``value()``."""
    return _Streaming("iou", reference_compat)


def mean_accuracy():
    return _Streaming("acc")
