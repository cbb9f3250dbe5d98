"""ResNet generator, the reference's DeepLab-style segmentation network, and canonical presets.

Re-design of the reference's ``core/resnet.py``:

* ``Bottleneck`` / ``BasicBlock``     -- core/resnet.py:94-152 / :57-91 (pre-activation units with the
  reference's post-add ReLU; stride applied after the 3x3 conv -- computed here directly at the
  strided output positions, same math, a quarter of the FLOPs, SURVEY K7 / defect 13)
* ``stack_blocks_plan``               -- resnet_utils.stack_blocks_dense bookkeeping (A7)
* ``ResNetV2Beta``                    -- core/resnet.py:171-257 (root, pool1, postnorm, blocks, pool5, logits)
* ``resnet_v2``                       -- core/resnet.py:284-354 (depths 128/258/512 + multi-grid block4)
* ``ResNetSegModel`` (= resnet_model) -- core/resnet.py:398-496 (backbone + "assp" + decoder)
* ``resnet50/101/152``                -- canonical v1.5 classification networks (BASELINE.json configs),
  built by the same generator with a different unit.

Activations are NHWC.  ``model_name`` prefixes every TF variable name and end-point key exactly like
``tf.variable_scope(model_name)`` does in model.py:353.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops import layers as L
from ..ops.layers import BNConfig, BatchNorm, Conv2d, ConvBNAct, SplitSeparableConv2d, TFModule

_DEFAULT_MULTI_GRID = [2, 2, 2]  # reference: core/resnet.py multi_grid default


class EndPoints(dict):
    """End-point collection (core/resnet.py:171-257 ``end_points_collection``).  ``wanted`` names the pre-activation
    tensors a caller will read; units whose ``.../conv3`` output is not wanted fuse the shortcut add and the ReLU
    into that convolution and never materialise it."""

    def __init__(self, wanted=()):
        super().__init__()
        self.wanted = frozenset(wanted)


# ---------------------------------------------------------------------------------------------
# units
# ---------------------------------------------------------------------------------------------
class Bottleneck(TFModule):
    """Reference bottleneck (core/resnet.py:94-152): preact BN+ReLU; 1x1 -> 3x3(dilated) -> [stride] -> 1x1;
    shortcut = subsample(inputs) if depth == depth_in else 1x1 conv(preact); relu(shortcut + residual)."""

    def __init__(self, depth_in: int, depth: int, depth_bottleneck: int, stride: int, rate: int, scope: str,
                 bn: BNConfig):
        super().__init__()
        s = f"{scope}/bottleneck_v2"
        self.scope, self.stride, self.depth_in, self.depth = s, stride, depth_in, depth
        self.preact = BatchNorm(depth_in, bn.decay, bn.eps, bn.scale, scope=f"{s}/preact")
        self.shortcut = (None if depth == depth_in else
                         Conv2d(depth_in, depth, 1, stride, use_bias=True, scope=f"{s}/shortcut",
                                weight_decay=bn.weight_decay))
        # The bottleneck width may not be a multiple of 8 (the reference's 258): it is stored channel-padded inside
        # the unit (zeros), which keeps every tensor 16-byte aligned for TMA without touching checkpoint shapes.
        self.conv1 = ConvBNAct(depth_in, depth_bottleneck, 1, scope=f"{s}/conv1", bn=bn, pad_out=True)
        # the reference runs this 3x3 at full resolution and then keeps every stride-th pixel; a strided conv
        # whose padding is the stride-1 SAME padding produces exactly those pixels.
        self.conv2 = ConvBNAct(depth_bottleneck, depth_bottleneck, 3, stride=stride, rate=rate,
                               padding=(rate, rate, rate, rate), scope=f"{s}/Conv", bn=bn, pad_in=True, pad_out=True)
        self.conv3 = Conv2d(depth_bottleneck, depth, 1, use_bias=True, scope=f"{s}/conv3",
                            weight_decay=bn.weight_decay, pad_in=True)

    def forward(self, x, end_points: Optional[dict] = None, prefix: str = ""):
        # The pre-activation BatchNorm hands out an alias ("tap") of its input: whatever else reads the unit input
        # (identity shortcut, subsampled shortcut) reads the alias, so that gradient comes back to the BatchNorm node
        # and is added -- and the input's own ReLU backward applied -- inside its backward kernel.  A projection
        # shortcut reads the pre-activation instead, through an alias handed out by conv1: its gradient is added on the
        # tensor core inside conv1's data-gradient kernel.  No aten::add_ / threshold_backward per unit either way.
        pre, xa = self.preact(x, relu=True, want_tap=True)
        if self.shortcut is None:
            sc = xa if self.stride == 1 else L.subsample(xa, self.stride)
            r = self.conv1(pre)
        elif self.training and torch.is_grad_enabled():
            r, pa = self.conv1(pre, want_tap=True)
            sc = self.shortcut(pa)
        else:
            r = self.conv1(pre)
            sc = self.shortcut(pre)
        r = self.conv2(r)
        want_r = end_points is not None and f"{prefix}{self.scope}/conv3" in getattr(end_points, "wanted", ())
        if end_points is None or (hasattr(end_points, "wanted") and not want_r):
            # relu(shortcut + conv3(.)) inside conv3's epilogue: no separate add / ReLU kernels
            # (the statistics of that sum for the next unit's pre-activation BatchNorm could ride along --
            # conv2d(..., want_stats=True, residual=...) -- but measured slower than the separate statistics pass:
            # 6.92 vs 6.82 ms/step, the epilogue of these small-map layers is the critical path)
            out = self.conv3(r, residual=sc, relu=True)
        else:
            r = self.conv3(r)
            out = F.relu(sc + r)
            end_points[f"{prefix}{self.scope}/conv3"] = r
        if end_points is not None:
            end_points[f"{prefix}{self.scope}"] = out
        return out


class BasicBlock(TFModule):
    """Reference basic_block (core/resnet.py:57-91): output depth is ``depth_bottleneck``."""

    def __init__(self, depth_in: int, depth: int, depth_bottleneck: int, stride: int, rate: int, scope: str,
                 bn: BNConfig):
        super().__init__()
        s = f"{scope}/bottleneck_v2"
        self.scope, self.stride = s, stride
        self.preact = BatchNorm(depth_in, bn.decay, bn.eps, bn.scale, scope=f"{s}/preact")
        self.shortcut = (None if depth_bottleneck == depth_in else
                         Conv2d(depth_in, depth_bottleneck, 1, stride, use_bias=True, scope=f"{s}/shortcut",
                                weight_decay=bn.weight_decay))
        self.conv1 = ConvBNAct(depth_in, depth_bottleneck, 3, stride=stride, padding=(1, 1, 1, 1), scope=f"{s}/Conv",
                               bn=bn)
        self.conv2 = Conv2d(depth_bottleneck, depth_bottleneck, 3, rate=rate, use_bias=True, scope=f"{s}/conv2",
                            weight_decay=bn.weight_decay)

    def forward(self, x, end_points: Optional[dict] = None, prefix: str = ""):
        pre = self.preact(x, relu=True)
        sc = L.subsample(x, self.stride) if self.shortcut is None else self.shortcut(pre)
        r = self.conv2(self.conv1(pre))
        out = F.relu(sc + r)
        if end_points is not None:
            end_points[f"{prefix}{self.scope}/conv2"] = r
            end_points[f"{prefix}{self.scope}"] = out
        return out


class BottleneckV1(TFModule):
    """Canonical ResNet v1.5 bottleneck: conv-BN-ReLU x2, conv-BN, projection shortcut with BN, stride on
    the 3x3, residual add + ReLU fused into the last BatchNorm kernel."""

    def __init__(self, depth_in: int, depth: int, depth_bottleneck: int, stride: int, rate: int, scope: str,
                 bn: BNConfig):
        super().__init__()
        s = f"{scope}/bottleneck_v1"
        self.scope = s
        self.conv1 = ConvBNAct(depth_in, depth_bottleneck, 1, scope=f"{s}/conv1", bn=bn)
        self.conv2 = ConvBNAct(depth_bottleneck, depth_bottleneck, 3, stride=stride, rate=rate,
                               padding=(rate, rate, rate, rate), scope=f"{s}/conv2", bn=bn)
        self.conv3 = ConvBNAct(depth_bottleneck, depth, 1, scope=f"{s}/conv3", bn=bn, relu=True)
        self.shortcut = (None if (depth == depth_in and stride == 1) else
                         ConvBNAct(depth_in, depth, 1, stride=stride, scope=f"{s}/shortcut", bn=bn, relu=False))

    def forward(self, x, end_points: Optional[dict] = None, prefix: str = ""):
        if self.training and torch.is_grad_enabled():
            # the shortcut branch (identity or projection) starts from conv1's input alias, so that in backward its
            # gradient reaches conv1's node and is added inside conv1's data-gradient kernel (on the tensor core,
            # csrc/conv_tcgen05.cu kResMma) instead of by a separate elementwise add
            h, xa = self.conv1(x, want_tap=True)
            sc = xa if self.shortcut is None else self.shortcut(xa)
        else:
            h = self.conv1(x)
            sc = x if self.shortcut is None else self.shortcut(x)
        out = self.conv3(self.conv2(h), residual=sc)  # relu(bn(conv3) + shortcut), one kernel
        if end_points is not None:
            end_points[f"{prefix}{self.scope}"] = out
        return out


UNIT_FNS = {"bottleneck": Bottleneck, "basic_block": BasicBlock, "bottleneck_v1": BottleneckV1}


# ---------------------------------------------------------------------------------------------
# block bookkeeping
# ---------------------------------------------------------------------------------------------
@dataclass
class Block:
    """resnet_utils.Block: scope, unit constructor name, and one arg dict per unit."""
    scope: str
    unit_fn: str
    args: List[dict] = field(default_factory=list)


def resnet_v2_beta_block(scope: str, base_depth: int, num_units: int, stride: int, block_fn: str,
                         stride_first: bool = False) -> Block:
    """core/resnet.py:260-281: (n-1) stride-1 units then one unit carrying the stride (last, like the
    reference; ``stride_first`` puts it on the first unit for the canonical v1.5 presets)."""
    unit = {"depth": base_depth * 4, "depth_bottleneck": base_depth, "stride": 1, "unit_rate": 1}
    args = [dict(unit) for _ in range(num_units)]
    args[0 if stride_first else -1]["stride"] = stride
    return Block(scope, block_fn, args)


def stack_blocks_plan(blocks: Sequence[Block], output_stride: Optional[int]) -> List[List[Tuple[int, int]]]:
    """resnet_utils.stack_blocks_dense bookkeeping (A7): per unit (stride, rate) after atrous conversion."""
    current_stride, rate = 1, 1
    plan = []
    for block in blocks:
        units = []
        for unit in block.args:
            if output_stride is not None and current_stride > output_stride:
                raise ValueError("The target output_stride cannot be reached.")
            s = unit.get("stride", 1)
            if output_stride is not None and current_stride == output_stride:
                units.append((1, rate * unit.get("unit_rate", 1)))
                rate *= s
            else:
                units.append((s, unit.get("unit_rate", 1)))
                current_stride *= s
        plan.append(units)
    if output_stride is not None and current_stride != output_stride:
        raise ValueError("The target output_stride cannot be reached.")
    return plan


# ---------------------------------------------------------------------------------------------
# backbone
# ---------------------------------------------------------------------------------------------
class RootBeta(TFModule):
    """root_block_fn_for_beta_variant (core/resnet.py:155-168): 3x3 s2 64, 3x3 64, 3x3 128."""

    def __init__(self, cin: int, scope: str, bn: BNConfig):
        super().__init__()
        self.conv1_1 = ConvBNAct(cin, 64, 3, stride=2, scope=f"{scope}/conv1_1", bn=bn)
        self.conv1_2 = ConvBNAct(64, 64, 3, scope=f"{scope}/conv1_2", bn=bn)
        self.conv1_3 = ConvBNAct(64, 128, 3, scope=f"{scope}/conv1_3", bn=bn)
        self.out_depth = 128

    def forward(self, x):
        return self.conv1_3(self.conv1_2(self.conv1_1(x)))


class Root7x7(TFModule):
    """Default root: conv2d_same 7x7 s2 64 (core/resnet.py:219-224); torchvision padding when symmetric."""

    def __init__(self, cin: int, scope: str, bn: BNConfig, symmetric: bool = False):
        super().__init__()
        self.conv1 = ConvBNAct(cin, 64, 7, stride=2, padding=(3, 3, 3, 3) if symmetric else "SAME",
                               scope=f"{scope}/conv1", bn=bn)
        self.out_depth = 64

    def forward(self, x):
        return self.conv1(x)


class ResNetV2Beta(TFModule):
    """Generator (core/resnet.py:171-257).  ``forward`` returns (net, end_points)."""

    def __init__(self, in_channels: int, blocks: Sequence[Block], num_classes: Optional[int] = None,
                 global_pool: bool = True, output_stride: Optional[int] = None, root: str = "beta",
                 scope: str = "resnet_v2", bn: Optional[BNConfig] = None, postnorm: bool = True,
                 symmetric_padding: bool = False, model_name: str = ""):
        super().__init__()
        bn = bn or BNConfig()
        self.scope = scope
        self.prefix = f"{model_name}/" if model_name else ""
        self.global_pool, self.num_classes = global_pool, num_classes
        self.symmetric_padding = symmetric_padding
        if output_stride is not None:
            if output_stride % 4 != 0:
                raise ValueError("The output_stride needs to be a multiple of 4.")
            output_stride //= 4  # (defect 12 fixed: only divided when given)
        self.root = RootBeta(in_channels, scope, bn) if root == "beta" else Root7x7(in_channels, scope, bn,
                                                                                    symmetric_padding)
        depth = self.root.out_depth
        self.postnorm = BatchNorm(depth, bn.decay, bn.eps, bn.scale, scope=f"{scope}/postnorm") if postnorm else None
        plan = stack_blocks_plan(blocks, output_stride)
        self.block_scopes: List[str] = []
        self.blocks = nn.ModuleList()
        self.unit_plan = plan
        for block, units in zip(blocks, plan):
            mods = nn.ModuleList()
            fn = UNIT_FNS[block.unit_fn]
            for i, (unit, (stride, rate)) in enumerate(zip(block.args, units)):
                m = fn(depth, unit["depth"], unit["depth_bottleneck"], stride, rate,
                       scope=f"{scope}/{block.scope}/unit_{i + 1}", bn=bn)
                depth = unit["depth_bottleneck"] if block.unit_fn == "basic_block" else unit["depth"]
                mods.append(m)
            self.blocks.append(mods)
            self.block_scopes.append(f"{scope}/{block.scope}")
        self.out_depth = depth
        self.logits = (Conv2d(depth, num_classes, 1, use_bias=True, scope=f"{scope}/logits",
                              weight_decay=bn.weight_decay) if num_classes is not None else None)

    def forward(self, x, collect: bool = True, wanted=None):
        """``wanted`` (optional): the only pre-activation end points the caller needs; see :class:`EndPoints`."""
        end_points: Optional[Dict[str, torch.Tensor]] = None
        if collect:
            end_points = {} if wanted is None else EndPoints(wanted)
        net = self.root(x)
        if self.symmetric_padding:
            from ..ops import pooling as P
            net = P.max_pool(net, 3, 2, (1, 1, 1, 1))
            pool_stats = None
        elif self.postnorm is not None and self.training:
            # pool1 -> postnorm (core/resnet.py:241-242): the pooling kernel leaves the BatchNorm statistics of its
            # output, the BatchNorm runs no statistics pass of its own
            net, pool_stats = L.max_pool_same(net, 3, 2, want_stats=True)
        else:
            net, pool_stats = L.max_pool_same(net, 3, 2), None
        if self.postnorm is not None:
            net = self.postnorm(net, relu=True, stats=pool_stats)
        for mods, bscope in zip(self.blocks, self.block_scopes):
            for m in mods:
                net = m(net, end_points, self.prefix)
            if collect:
                end_points[f"{self.prefix}{bscope}"] = net
        if self.global_pool:
            net = L.global_avg_pool(net, keepdims=True)
            if collect:
                end_points["global_pool"] = net
        if self.logits is not None:
            net = self.logits(net)
            if collect:
                end_points[f"{self.prefix}{self.scope}/logits"] = net
                end_points["predictions"] = torch.softmax(net.float(), dim=-1)
        return net, end_points


def resnet_v2(in_channels: int, n_blocks=(3, 4, 6), block_type: str = "bottleneck", num_classes=None,
              global_pool: bool = False, output_stride: Optional[int] = None, multi_grid=None,
              scope: str = "resnet_v2_34", bn: Optional[BNConfig] = None, model_name: str = "") -> ResNetV2Beta:
    """The reference's network builder (core/resnet.py:284-354), depths 128 / 258 (sic) / 512 + multi-grid block4."""
    if multi_grid is None:
        multi_grid = _DEFAULT_MULTI_GRID
    elif len(multi_grid) != 3:
        raise ValueError("Expect multi_grid to have length 3.")
    if block_type not in ("bottleneck", "basic_block"):
        block_type = "bottleneck"
    blocks = [
        resnet_v2_beta_block("block1", 128, n_blocks[0], 2, block_type),
        resnet_v2_beta_block("block2", 258, n_blocks[1], 2, block_type),
        resnet_v2_beta_block("block3", 512, n_blocks[2], 2, block_type),
        Block("block4", block_type,
              [{"depth": 1024, "depth_bottleneck": 256, "stride": 1, "unit_rate": r} for r in multi_grid]),
    ]
    return ResNetV2Beta(in_channels, blocks, num_classes, global_pool, output_stride, root="beta", scope=scope, bn=bn,
                        model_name=model_name)


# ---------------------------------------------------------------------------------------------
# canonical classification presets
# ---------------------------------------------------------------------------------------------
class ResNetClassifier(TFModule):
    """ResNet v1.5 (50/101/152): 7x7 s2 root, 3x3 s2 max-pool, 4 stages, global pool, FC as a 1x1 conv."""

    def __init__(self, layers: Sequence[int], num_classes: int = 1000, in_channels: int = 3,
                 bn: Optional[BNConfig] = None, scope: str = "resnet_v1"):
        super().__init__()
        bn = bn or BNConfig(decay=0.9, eps=1e-5)
        blocks = [resnet_v2_beta_block(f"block{i + 1}", 64 << i, n, 1 if i == 0 else 2, "bottleneck_v1",
                                       stride_first=True) for i, n in enumerate(layers)]
        self.net = ResNetV2Beta(in_channels, blocks, num_classes=num_classes, global_pool=True, output_stride=None,
                                root="7x7", scope=scope, bn=bn, postnorm=False, symmetric_padding=True)
        # The 3-channel image goes to the first layer unpadded: the stem kernel packs it space-to-depth itself
        # (csrc/conv_tcgen05.cu, kStem).  The trainer reads this attribute; a value would make it zero-pad the
        # channels on the device (the generic kernel on a padded image measured 5.3 ms against the stem kernel's 0.18).
        self.input_channel_pad = None
        for m in self.modules():  # zero-init the last BN gamma of every unit (standard large-batch recipe)
            if isinstance(m, BottleneckV1) and m.conv3.bn.gamma is not None:
                nn.init.zeros_(m.conv3.bn.gamma)

    def forward(self, x):
        net, _ = self.net(x, collect=False)
        return net.reshape(net.shape[0], -1)


def resnet50(num_classes: int = 1000, **kw) -> ResNetClassifier:
    return ResNetClassifier((3, 4, 6, 3), num_classes, scope="resnet_v1_50", **kw)


def resnet101(num_classes: int = 1000, **kw) -> ResNetClassifier:
    return ResNetClassifier((3, 4, 23, 3), num_classes, scope="resnet_v1_101", **kw)


def resnet152(num_classes: int = 1000, **kw) -> ResNetClassifier:
    return ResNetClassifier((3, 8, 36, 3), num_classes, scope="resnet_v1_152", **kw)


class SmallResNet(TFModule):
    """BASELINE config 1: a small classifier from the same generator (32x32x3, 10 classes) for CPU-tower tests."""

    def __init__(self, num_classes: int = 10, in_channels: int = 3, width: int = 16, bn: Optional[BNConfig] = None):
        super().__init__()
        bn = bn or BNConfig(decay=0.9, eps=1e-5)
        blocks = [resnet_v2_beta_block("block1", width, 1, 2, "bottleneck"),
                  resnet_v2_beta_block("block2", 2 * width, 1, 1, "bottleneck")]
        self.net = ResNetV2Beta(in_channels, blocks, num_classes=num_classes, global_pool=True, root="beta",
                                scope="resnet_small", bn=bn)

    def forward(self, x):
        net, _ = self.net(x, collect=False)
        return net.reshape(net.shape[0], -1)


# ---------------------------------------------------------------------------------------------
# the reference's full segmentation network
# ---------------------------------------------------------------------------------------------
class ResNetSegModel(TFModule):
    """``resnet_model`` (core/resnet.py:398-496): backbone -> "assp" -> upsample -> decoder -> logits.

    ``forward(images[N,H,W,C]) -> logits[N,H,W,1]``.  The reference hard-codes the (26, 26) decoder size for
    101x101 inputs (core/resnet.py:474, defect 11); here it is taken from the low-level feature map, which
    is the same number for the default input.
    """

    def __init__(self, model_name: str = "model", in_channels: int = 2, weight_decay: float = 0.001,
                 batch_norm_decay: float = 0.99, batch_norm_epsilon: float = 0.001, batch_norm_scale: bool = True,
                 output_stride: int = 8, base_depth: int = 256, input_shape=(101, 101), n_blocks=(3, 4, 6),
                 block_type: str = "bottleneck"):
        super().__init__()
        if len(n_blocks) != 3:
            raise ValueError("Expect n_blocks to have length 3.")
        bn = BNConfig(batch_norm_decay, batch_norm_epsilon, batch_norm_scale, weight_decay)
        self.model_name, self.input_shape, self.block_type = model_name, tuple(input_shape), block_type
        P = f"{model_name}/"
        self.backbone = resnet_v2(in_channels, n_blocks, block_type, None, False, output_stride, (1, 2, 1),
                                  scope=f"{P}resnet_v2", bn=bn)
        c4 = self.backbone.out_depth
        self.aspp_1x1 = ConvBNAct(c4, base_depth, 1, scope=f"{P}assp/conv/conv_1x1", bn=bn)
        self.aspp_sep = nn.ModuleList([
            SplitSeparableConv2d(c4, base_depth, 3, rate, scope=f"{P}assp/conv/conv_3x3_{i + 1}", bn=bn)
            for i, rate in enumerate((2, 4, 8))])
        self.aspp_pool = ConvBNAct(c4, base_depth, 1, scope=f"{P}assp/pooling/conv_1x1", bn=bn)
        self.aspp_out = ConvBNAct(5 * base_depth, base_depth, 1, scope=f"{P}assp/conv_1x1", bn=bn)
        low = self.backbone.blocks[0][0]
        c_low = low.conv2.cout if block_type == "basic_block" else low.depth
        self.low_key = f"{low.scope}/conv{2 if block_type == 'basic_block' else 3}"
        self.block4_key = f"{P}resnet_v2/block4"
        self.dec_1x1 = ConvBNAct(c_low, base_depth, 1, scope=f"{P}decoder/conv_1x1", bn=bn)
        self.dec_3x3 = Conv2d(2 * base_depth, 1, 3, use_bias=True, scope=f"{P}decoder/conv_3x3",
                              weight_decay=weight_decay)

    @staticmethod
    def _branch(m: ConvBNAct, x, upsample: bool = False):
        """conv of a ConvBNAct, with its BatchNorm left for the concat kernel to apply (straight into a channel slice)."""
        if m.training:
            y, stats = m.conv(x, want_stats=True)
        else:
            y, stats = m.conv(x), None
        return {"x": y, "bn": m.bn, "relu": m.relu, "stats": stats, "upsample": upsample}

    def forward(self, x):
        from ..ops import functional as TF

        _, ep = self.backbone(x, collect=True, wanted=(self.low_key,))
        feat = ep[self.block4_key]
        size = feat.shape[1:3]
        low = ep[self.low_key]
        if TF._native_ok(feat) and all(m.bn.c_p % 8 == 0 for m in (self.aspp_1x1, self.aspp_pool, self.dec_1x1)):
            from ..ops import pooling as PL

            # K10: image pooling = global mean folded into its 1x1 convolution (core/resnet.py:464-466), one kernel
            pooled_y = PL.gap_conv1x1(feat, self.aspp_pool.conv.weight)
            # K12: no concatenation -- every branch's BatchNorm (the pooled branch: BatchNorm + `_upsample`) writes
            # its channel slice of ONE buffer (core/resnet.py:470); same for the decoder's pair (core/resnet.py:485)
            aspp_in = TF.batch_norm_concat(
                [self._branch(self.aspp_1x1, feat)] +
                [self._branch(m.pointwise, m.depthwise(feat)) for m in self.aspp_sep] +
                [{"x": pooled_y, "bn": self.aspp_pool.bn, "relu": self.aspp_pool.relu, "stats": None,
                  "upsample": True}], size)
            aspp = self.aspp_out(aspp_in)
            dec = TF.batch_norm_concat([self._branch(self.dec_1x1, low), {"x": aspp, "bn": None, "upsample": True}],
                                       low.shape[1:3])
        else:
            pooled_in = L.global_avg_pool(feat, keepdims=True)
            branches = [self.aspp_1x1(feat)] + [m(feat) for m in self.aspp_sep]
            branches.append(L.upsample(self.aspp_pool(pooled_in), size))
            aspp = self.aspp_out(torch.cat(branches, dim=-1))
            dec = torch.cat([self.dec_1x1(low), L.upsample(aspp, low.shape[1:3])], dim=-1)
        dec = self.dec_3x3(dec)
        return L.upsample(dec.float(), self.input_shape)


def resnet_model(**kw) -> ResNetSegModel:
    return ResNetSegModel(**kw)
