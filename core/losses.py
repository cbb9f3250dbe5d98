"""Reference-compatible ``core.losses`` (reference: core/losses.py:5-92), the reference's signatures incl. ``ignore``."""
from tensorflowdistributedlearning_b200.compat import flatten_binary_scores, lovasz_hinge, lovasz_loss  # noqa: F401
from tensorflowdistributedlearning_b200.ops.reference import lovasz_grad, lovasz_hinge_flat  # noqa: F401
