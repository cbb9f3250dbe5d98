"""Xception-41 backbone (the *intended* network of the reference's dormant ``core/xception.py``) plus a
classification head.

The reference file does not import (``resnet_utils`` undefined, core/xception.py:352), builds only the
last unit of each block (indent bug, :272-284) and has no head (defect 15).  This module
implements what the code evidently means:

* ``fixed_padding``             -- core/xception.py:18-35
* ``SeparableConv2dSame``       -- :38-128 (depthwise -> pointwise, explicit padding + VALID when strided)
* ``XceptionModule``            -- :131-228 (3 separable convs, stride on the third, conv / sum / none skip)
* ``stack plan``                -- :231-292 (atrous bookkeeping, all units built)
* ``Xception`` / ``xception_41``-- :295-364 / :405-465 (entry 128-256-728, 8 middle units, exit 1024-2048)
* head (new)                    -- global average pool -> dropout(keep_prob) -> 1x1 logits

BatchNorm follows every depthwise and pointwise conv (DeepLab's xception arg scope, which the reference
file relies on but does not define).
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import os

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops import layers as L
from ..ops.layers import BNConfig, BatchNorm, Conv2d, ConvBNAct, DepthwiseConv2d, TFModule


def fixed_padding(x, kernel_size: int, rate: int = 1):
    return L.fixed_padding(x, kernel_size, rate)


def _explicit_pads(kernel_size: int, rate: int):
    k_eff = kernel_size + (kernel_size - 1) * (rate - 1)
    total = k_eff - 1
    beg = total // 2
    return (beg, total - beg, beg, total - beg)


class SeparableConv2dSame(TFModule):
    """depthwise kxk (+BN, optional ReLU) -> pointwise 1x1 (+BN, optional ReLU); 'SAME' output size with
    explicit symmetric-first padding when strided (core/xception.py:38-128)."""

    def __init__(self, cin: int, cout: int, kernel_size: int, stride: int, rate: int, scope: str, bn: BNConfig,
                 relu: bool, relu_in: bool = False):
        super().__init__()
        padding = "SAME" if stride == 1 else _explicit_pads(kernel_size, rate)
        self.depthwise = DepthwiseConv2d(cin, kernel_size, stride, rate, padding, use_bias=False, relu=False,
                                         scope=f"{scope}_depthwise", init_std=0.33, relu_in=relu_in)
        self.dw_bn = BatchNorm(cin, bn.decay, bn.eps, bn.scale, scope=f"{scope}_depthwise/BatchNorm")
        self.pointwise = ConvBNAct(cin, cout, 1, scope=f"{scope}_pointwise", bn=bn, relu=relu, init_std=0.06)
        self.relu = relu

    def forward(self, x, residual=None, want_tap: bool = False):
        """``residual`` (the module's skip branch) is added inside the pointwise BatchNorm kernel.
        ``want_tap=True`` (training): returns (out, x_alias) -- the module's skip connection reads ``x_alias`` so that
        its gradient is added inside the depthwise data-gradient kernel."""
        if want_tap:
            y, stats, tap = self.depthwise(x, want_stats=True, want_tap=True)
            return self.pointwise(self.dw_bn(y, relu=self.relu, stats=stats, defer_apply=0 if self.relu else 2),
                                  residual=residual), tap
        if self.training:  # the depthwise kernel accumulates the statistics of its own output for dw_bn
            y, stats = self.depthwise(x, want_stats=True)
            # no activation between this BatchNorm and the pointwise conv: the normalise pass is folded into the
            # pointwise weights instead of being run over the activation (ops/functional.batch_norm(defer_apply));
            # 2 = the pointwise conv itself feeds a training-mode BatchNorm (its output gradient has zero column sums)
            return self.pointwise(self.dw_bn(y, relu=self.relu, stats=stats, defer_apply=0 if self.relu else 2),
                                  residual=residual)
        return self.pointwise(self.dw_bn(self.depthwise(x), relu=self.relu), residual=residual)


class XceptionModule(TFModule):
    def __init__(self, cin: int, depth_list: Sequence[int], skip: str, stride: int, unit_rate_list: Sequence[int],
                 rate: int, act_in_sep: bool, scope: str, bn: BNConfig):
        super().__init__()
        if len(depth_list) != 3 or len(unit_rate_list) != 3:
            raise ValueError("Expect three elements in depth_list / unit_rate_list.")
        if skip not in ("conv", "sum", "none"):
            raise ValueError("Unsupported skip connection type.")
        s = f"{scope}/xception_module"
        self.scope, self.skip, self.act_in_sep = s, skip, act_in_sep
        convs, c = [], cin
        for i in range(3):
            # without activation inside the separable conv, the module applies ReLU *before* each one
            # (core/xception.py:184-210); that pre-activation is folded into the depthwise kernel's load
            convs.append(SeparableConv2dSame(c, depth_list[i], 3, stride if i == 2 else 1, rate * unit_rate_list[i],
                                             scope=f"{s}/separable_conv{i + 1}", bn=bn, relu=act_in_sep,
                                             relu_in=not act_in_sep))
            c = depth_list[i]
        self.convs = nn.ModuleList(convs)
        self.shortcut = (ConvBNAct(cin, depth_list[-1], 1, stride=stride, scope=f"{s}/shortcut", bn=bn, relu=False)
                         if skip == "conv" else None)
        self.out_depth = depth_list[-1]

    def forward(self, x):
        # the skip connection (core/xception.py:211-222) rides into the last pointwise conv's BatchNorm kernel as its
        # residual input instead of a separate elementwise add (that conv has no activation when there is a skip...
        if (self.skip != "none" and not self.act_in_sep and self.training and torch.is_grad_enabled()
                and os.environ.get("TDL_DW_TAP", "1") != "0"):
            # the skip branch reads an alias of x handed out by the first depthwise conv, so that branch's gradient is
            # added inside the depthwise data-gradient kernel (no aten::add per module in backward)
            r, xa = self.convs[0](x, want_tap=True)
            skip = self.shortcut(xa) if self.skip == "conv" else xa
            r = self.convs[1](r)
            return self.convs[2](r, residual=skip)
        skip = self.shortcut(x) if self.skip == "conv" else (x if self.skip == "sum" else None)
        if skip is not None and self.act_in_sep:
            # ...unless the unit applies ReLU inside its separable convs: then the sum must stay outside the activation
            r = x
            for m in self.convs:
                r = m(r)
            return r + skip
        r = self.convs[0](x)
        r = self.convs[1](r)
        return self.convs[2](r, residual=skip)


def xception_block(scope, depth_list, skip_connection_type, activation_fn_in_separable_conv, num_units, stride,
                   unit_rate_list=None):
    if unit_rate_list is None:
        unit_rate_list = [1, 1, 1]
    return {"scope": scope, "units": [{"depth_list": list(depth_list), "skip": skip_connection_type,
                                       "act_in_sep": activation_fn_in_separable_conv, "stride": stride,
                                       "unit_rate_list": list(unit_rate_list)} for _ in range(num_units)]}


class Xception(TFModule):
    """Generator (core/xception.py:295-364) + optional classification head."""

    def __init__(self, in_channels: int, blocks: List[dict], num_classes: Optional[int] = None, keep_prob: float = 0.5,
                 output_stride: Optional[int] = None, scope: str = "xception_41", bn: Optional[BNConfig] = None):
        super().__init__()
        bn = bn or BNConfig(decay=0.9997, eps=1e-3, scale=True, weight_decay=4e-5)
        self.scope = scope
        if output_stride is not None:
            if output_stride % 2 != 0:
                raise ValueError("The output_stride needs to be a multiple of 2.")
            output_stride //= 2
        self.conv1_1 = ConvBNAct(in_channels, 32, 3, stride=2, padding=_explicit_pads(3, 1),
                                 scope=f"{scope}/entry_flow/conv1_1", bn=bn)
        self.conv1_2 = ConvBNAct(32, 64, 3, scope=f"{scope}/entry_flow/conv1_2", bn=bn)
        current_stride, rate, c = 1, 1, 64
        units = []
        for block in blocks:
            for i, u in enumerate(block["units"]):
                if output_stride is not None and current_stride > output_stride:
                    raise ValueError("The target output_stride cannot be reached.")
                if output_stride is not None and current_stride == output_stride:
                    stride, r = 1, rate
                    rate *= u["stride"]
                else:
                    stride, r = u["stride"], 1
                    current_stride *= u["stride"]
                m = XceptionModule(c, u["depth_list"], u["skip"], stride, u["unit_rate_list"], r, u["act_in_sep"],
                                   scope=f"{scope}/{block['scope']}/unit_{i + 1}", bn=bn)
                c = m.out_depth
                units.append(m)
        if output_stride is not None and current_stride != output_stride:
            raise ValueError("The target output_stride cannot be reached.")
        self.units = nn.ModuleList(units)
        self.out_depth = c
        self.keep_prob = keep_prob
        self.logits = (Conv2d(c, num_classes, 1, use_bias=True, scope=f"{scope}/logits", weight_decay=bn.weight_decay)
                       if num_classes is not None else None)

    def features(self, x):
        net = self.conv1_2(self.conv1_1(x))
        for m in self.units:
            net = m(net)
        return net

    def forward(self, x):
        net = self.features(x)
        if self.logits is None:
            return net
        net = L.global_avg_pool(net, keepdims=True)
        if self.training and self.keep_prob < 1.0:
            net = F.dropout(net, 1.0 - self.keep_prob)
        return self.logits(net).reshape(net.shape[0], -1)


def xception_41(in_channels: int = 3, num_classes: Optional[int] = None, keep_prob: float = 0.5,
                output_stride: Optional[int] = None, multi_grid: Optional[Sequence[int]] = None,
                scope: str = "xception_41", bn: Optional[BNConfig] = None) -> Xception:
    blocks = [
        xception_block("entry_flow/block1", [128, 128, 128], "conv", False, 1, 2),
        xception_block("entry_flow/block2", [256, 256, 256], "conv", False, 1, 2),
        xception_block("entry_flow/block3", [728, 728, 728], "conv", False, 1, 2),
        xception_block("middle_flow/block1", [728, 728, 728], "sum", False, 8, 1),
        xception_block("exit_flow/block1", [728, 1024, 1024], "conv", False, 1, 2),
        xception_block("exit_flow/block2", [1536, 1536, 2048], "none", True, 1, 1, unit_rate_list=multi_grid),
    ]
    return Xception(in_channels, blocks, num_classes, keep_prob, output_stride, scope, bn)
