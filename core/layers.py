"""Reference-compatible ``core.layers`` (reference: core/layers.py:6-109): functional forms with the reference's
arguments (variables are created on first use under ``scope``, see ``tensorflowdistributedlearning_b200.compat``);
the module class is exported as ``SplitSeparableConv2d``."""
from tensorflowdistributedlearning_b200.compat import _fixed_padding, _upsample, split_separable_conv2d  # noqa: F401
from tensorflowdistributedlearning_b200.ops.layers import SplitSeparableConv2d  # noqa: F401
