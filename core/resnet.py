"""Reference-compatible ``core.resnet`` (reference: core/resnet.py): ``bottleneck`` / ``basic_block`` /
``resnet_model`` are callable with the reference's arguments; the generator pieces are the package's builders."""
from tensorflowdistributedlearning_b200.compat import basic_block, bottleneck, resnet_model  # noqa: F401
from tensorflowdistributedlearning_b200.models.resnet import (BasicBlock, Bottleneck, ResNetSegModel,  # noqa: F401
                                                              ResNetV2Beta as resnet_v2_beta,
                                                              RootBeta as root_block_fn_for_beta_variant,
                                                              resnet_v2, resnet_v2_beta_block, stack_blocks_plan)
from tensorflowdistributedlearning_b200.ops.layers import BNConfig as resnet_arg_scope  # noqa: F401
