"""Reference-compatible ``core.losses`` (reference: core/losses.py)."""
from tensorflowdistributedlearning_b200.ops.functional import lovasz_loss as _lovasz_loss
from tensorflowdistributedlearning_b200.ops.reference import (lovasz_grad, lovasz_hinge_flat,  # noqa: F401
                                                              lovasz_hinge_per_image)


def lovasz_loss(y_true, y_pred, data_format="NHWC"):
    """Per-image Lovasz hinge of logits ``y_pred`` averaged over the batch; NCHW inputs are squeezed on axis 1."""
    if data_format == "NCHW":
        y_true, y_pred = y_true.squeeze(1), y_pred.squeeze(1)
    return _lovasz_loss(y_true, y_pred)


def lovasz_hinge(logits, labels, per_image=True, ignore=None):
    if ignore is not None:
        raise NotImplementedError("ignore labels are not used by the reference's caller")
    if per_image:
        return lovasz_hinge_per_image(logits.reshape(logits.shape[0], -1).float(),
                                      labels.reshape(labels.shape[0], -1).float()).mean()
    return lovasz_hinge_flat(logits.reshape(-1).float(), labels.reshape(-1).float())
