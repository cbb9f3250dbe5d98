"""Reference-compatible ``core.resnet`` (reference: core/resnet.py)."""
from tensorflowdistributedlearning_b200.models.resnet import (BasicBlock as basic_block, Bottleneck as bottleneck,  # noqa: F401
                                                              ResNetSegModel, ResNetV2Beta as resnet_v2_beta,
                                                              RootBeta as root_block_fn_for_beta_variant, resnet_model,
                                                              resnet_v2, resnet_v2_beta_block, stack_blocks_plan)
from tensorflowdistributedlearning_b200.ops.layers import BNConfig as resnet_arg_scope  # noqa: F401
