"""Reference-compatible ``core.layers`` (reference: core/layers.py): module-style equivalents of the slim helpers."""
from tensorflowdistributedlearning_b200.ops.layers import (SplitSeparableConv2d as split_separable_conv2d,  # noqa: F401
                                                           fixed_padding as _fixed_padding, upsample as _upsample)
