"""Reference-compatible ``core.xception`` (reference: core/xception.py, repaired)."""
from tensorflowdistributedlearning_b200.compat import split_separable_conv2d  # noqa: F401
from tensorflowdistributedlearning_b200.models.xception import (SeparableConv2dSame as separable_conv2d_same,  # noqa: F401
                                                                Xception as xception, XceptionModule as xception_module,
                                                                fixed_padding, xception_41, xception_block)
