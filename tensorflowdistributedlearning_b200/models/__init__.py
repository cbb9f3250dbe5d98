"""Model zoo: the reference's ResNet generator + DeepLab-style head, canonical ResNets, Xception-41."""
from .resnet import (BasicBlock, Block, Bottleneck, BottleneckV1, ResNetClassifier, ResNetSegModel, ResNetV2Beta,
                     SmallResNet, resnet50, resnet101, resnet152, resnet_model, resnet_v2, resnet_v2_beta_block,
                     stack_blocks_plan)
from .xception import Xception, XceptionModule, xception_41


def build(name: str, **kw):
    """Factory used by the CLI / bench: name in {resnet50, resnet101, resnet152, xception41, refseg, small}."""
    name = name.lower().replace("-", "").replace("_", "")
    if name == "resnet50":
        return resnet50(**kw)
    if name == "resnet101":
        return resnet101(**kw)
    if name == "resnet152":
        return resnet152(**kw)
    if name in ("xception", "xception41"):
        kw.setdefault("num_classes", 1000)
        return xception_41(**kw)
    if name in ("refseg", "resnetmodel", "seg"):
        return ResNetSegModel(**kw)
    if name in ("small", "smallresnet"):
        return SmallResNet(**kw)
    raise ValueError(f"unknown model {name}")
