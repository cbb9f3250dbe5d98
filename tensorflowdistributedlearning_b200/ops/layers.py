"""NHWC layer modules with TF-slim semantics and TF variable naming.

These replace the ``slim.conv2d`` / ``slim.batch_norm`` / ``slim.separable_conv2d`` / pooling layers
and the three helpers of the reference's ``core/layers.py`` (``split_separable_conv2d`` :6-49,
``_fixed_padding`` :52-79, ``_upsample`` :82-109).  Every module carries the *full TF variable
scope* it would have had under slim (e.g. ``resnet_v2/block1/unit_1/bottleneck_v2/conv1``), which
is what the checkpoint writer uses to keep the reference's variable layout loadable
(SURVEY.md section 2.7).

Layout: activations are [N, H, W, C]; conv weights are stored [Cout, kh, kw, Cin] (the layout the
tcgen05 kernels consume) and exported to checkpoints as TF's [kh, kw, Cin, Cout].
"""
from __future__ import annotations

import math
from typing import Dict, Iterator, List, Optional, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import functional as TF
from . import reference as R


class TFModule(nn.Module):
    """Module that knows the TF variable names of its own (non-recursive) tensors."""

    def tf_variables(self) -> Dict[str, Tuple[torch.Tensor, str]]:
        """name -> (tensor, kind); kind in {'conv', 'depthwise', 'plain'} controls the layout transform."""
        return {}


def _unpad(t: torch.Tensor, shape) -> torch.Tensor:
    """Leading sub-block of ``t`` with the logical (checkpoint) ``shape``; ``t`` itself when nothing is padded."""
    if shape is None or tuple(t.shape) == tuple(shape):
        return t
    return t[tuple(slice(0, n) for n in shape)]


def iter_tf_variables_ex(model: nn.Module):
    """(name, storage tensor, kind, logical shape).  Storage may be channel-padded to a multiple of 8 for the TMA /
    vector kernels (e.g. the reference's 258-channel block); the logical shape is what checkpoints hold."""
    for m in model.modules():
        if isinstance(m, TFModule):
            for name, v in m.tf_variables().items():
                t, kind = v[0], v[1]
                yield name, t, kind, (v[2] if len(v) > 2 else None)


def iter_tf_variables(model: nn.Module) -> Iterator[Tuple[str, torch.Tensor, str]]:
    """(name, tensor with its logical shape, kind) -- a view when the storage is channel-padded."""
    for name, t, kind, shape in iter_tf_variables_ex(model):
        yield name, _unpad(t, shape), kind


def pad8(c: int) -> int:
    return (c + 7) // 8 * 8


def to_tf_layout(t: torch.Tensor, kind: str) -> torch.Tensor:
    if kind == "conv":        # [K,R,S,C] -> [R,S,C,K]
        return t.permute(1, 2, 3, 0).contiguous()
    if kind == "depthwise":   # [R,S,C] -> [R,S,C,1]
        return t.unsqueeze(-1).contiguous()
    return t


def from_tf_layout(t: torch.Tensor, kind: str) -> torch.Tensor:
    if kind == "conv":
        return t.permute(3, 0, 1, 2).contiguous()
    if kind == "depthwise":
        return t.squeeze(-1).contiguous()
    return t


class Conv2d(TFModule):
    """slim.conv2d without normaliser/activation: weights (+ biases).  padding 'SAME' | 'VALID'."""

    def __init__(self, cin: int, cout: int, k: int, stride: int = 1, rate: int = 1, padding: str = "SAME",
                 use_bias: bool = False, scope: str = "Conv", init_std: Optional[float] = None,
                 weight_decay: float = 0.0, pad_in: bool = False, pad_out: bool = False):
        """``pad_in`` / ``pad_out`` store the weight with Cin / Cout rounded up to a multiple of 8 (zero filled), for
        producers / consumers that keep channel-padded activations.  Padded entries see zero activations and zero
        gradients, so they stay exactly zero; checkpoints hold the logical [cout, k, k, cin] block only."""
        super().__init__()
        self.cin, self.cout, self.k, self.stride, self.rate, self.padding = cin, cout, k, stride, rate, padding
        self.scope = scope
        self.weight_decay = weight_decay  # recorded like slim's regulariser; see ParamStore for its use
        self.cin_p = pad8(cin) if pad_in else cin
        self.cout_p = pad8(cout) if pad_out else cout
        w = torch.empty(cout, k, k, cin)
        if init_std is None:
            R.variance_scaling_(w)
        else:
            R.truncated_normal_(w, init_std)
        full = torch.zeros(self.cout_p, k, k, self.cin_p)
        full[:cout, :, :, :cin] = w
        self.weight = nn.Parameter(full)
        self.bias = nn.Parameter(torch.zeros(self.cout_p)) if use_bias else None

    def forward(self, x: torch.Tensor, want_stats: bool = False, want_tap: bool = False,
                residual: Optional[torch.Tensor] = None, relu: bool = False):
        return TF.conv2d(x, self.weight, self.bias, self.stride, self.rate, self.padding, want_stats, want_tap,
                         residual, relu)

    def tf_variables(self):
        d = {f"{self.scope}/weights": (self.weight, "conv", (self.cout, self.k, self.k, self.cin))}
        if self.bias is not None:
            d[f"{self.scope}/biases"] = (self.bias, "plain", (self.cout,))
        return d


class BatchNorm(TFModule):
    """slim.batch_norm (fused): beta, optional gamma, moving_mean, moving_variance; optional fused
    residual add and ReLU.  ``decay`` is TF's moving-average decay (reference default 0.99)."""

    def __init__(self, c: int, decay: float = 0.99, eps: float = 1e-3, scale: bool = True, scope: str = "BatchNorm",
                 pad: bool = False):
        super().__init__()
        self.c, self.decay, self.eps, self.scope = c, decay, eps, scope
        cp = pad8(c) if pad else c  # padded channels carry zeros in, zeros out (beta 0) and zero gradients
        self.c_p = cp
        self.beta = nn.Parameter(torch.zeros(cp))
        self.gamma = nn.Parameter(torch.ones(cp)) if scale else None
        self.register_buffer("moving_mean", torch.zeros(cp))
        self.register_buffer("moving_variance", torch.ones(cp))

    def forward(self, x: torch.Tensor, relu: bool = False, residual: Optional[torch.Tensor] = None,
                stats: Optional[torch.Tensor] = None, want_tap: bool = False, defer_apply: bool = False):
        if stats is None and self.training:
            # partial statistics the producing kernel left on the tensor (e.g. a unit tail relu(shortcut + conv3))
            stats = getattr(x, "_tdl_stats", None)
            if stats is not None and stats.shape[-1] != x.shape[-1]:
                stats = None
        return TF.batch_norm(x, self.gamma, self.beta, self.moving_mean, self.moving_variance, decay=self.decay,
                             eps=self.eps, training=self.training, relu=relu, residual=residual, stats=stats,
                             want_tap=want_tap, defer_apply=defer_apply)

    def tf_variables(self):
        d = {f"{self.scope}/beta": (self.beta, "plain", (self.c,)),
             f"{self.scope}/moving_mean": (self.moving_mean, "plain", (self.c,)),
             f"{self.scope}/moving_variance": (self.moving_variance, "plain", (self.c,))}
        if self.gamma is not None:
            d[f"{self.scope}/gamma"] = (self.gamma, "plain", (self.c,))
        return d


class BNConfig:
    """The subset of resnet_arg_scope (reference: core/resnet.py:357-395) that shapes layers."""

    def __init__(self, decay: float = 0.997, eps: float = 1e-5, scale: bool = True, weight_decay: float = 1e-4):
        self.decay, self.eps, self.scale, self.weight_decay = decay, eps, scale, weight_decay


class ConvBNAct(TFModule):
    """slim.conv2d under the resnet arg scope: conv (no bias) -> BatchNorm -> ReLU."""

    def __init__(self, cin, cout, k, stride=1, rate=1, padding="SAME", scope="Conv", bn: Optional[BNConfig] = None,
                 relu: bool = True, init_std: Optional[float] = None, pad_in: bool = False, pad_out: bool = False):
        super().__init__()
        bn = bn or BNConfig()
        self.conv = Conv2d(cin, cout, k, stride, rate, padding, use_bias=False, scope=scope, init_std=init_std,
                           weight_decay=bn.weight_decay, pad_in=pad_in, pad_out=pad_out)
        self.bn = BatchNorm(cout, bn.decay, bn.eps, bn.scale, scope=f"{scope}/BatchNorm", pad=pad_out)
        self.relu = relu

    def forward(self, x, residual: Optional[torch.Tensor] = None, want_tap: bool = False):
        """``want_tap=True`` -> (out, x_alias): use x_alias as the identity shortcut of the enclosing unit so that its
        gradient is added inside this convolution's data-gradient kernel."""
        if want_tap:
            y, stats, tap = self.conv(x, want_stats=self.training, want_tap=True)
            return self.bn(y, relu=self.relu, residual=residual, stats=stats), tap
        if self.training:  # the conv epilogue accumulates the BatchNorm statistics of its own output
            y, stats = self.conv(x, want_stats=True)
            return self.bn(y, relu=self.relu, residual=residual, stats=stats)
        return self.bn(self.conv(x), relu=self.relu, residual=residual)


class DepthwiseConv2d(TFModule):
    """slim.separable_conv2d(num_outputs=None): depthwise kxk (+bias) (+ReLU), depth multiplier 1."""

    def __init__(self, c: int, k: int = 3, stride: int = 1, rate: int = 1, padding="SAME", use_bias: bool = True,
                 relu: bool = True, scope: str = "depthwise", init_std: float = 0.33, relu_in: bool = False):
        super().__init__()
        self.c, self.k, self.stride, self.rate, self.padding, self.relu, self.scope = c, k, stride, rate, padding, relu, scope
        self.relu_in = relu_in  # apply ReLU to the input inside the kernel (Xception's pre-activation)
        self.weight = nn.Parameter(torch.empty(k, k, c))
        R.truncated_normal_(self.weight.data, init_std)
        self.bias = nn.Parameter(torch.zeros(c)) if use_bias else None

    def forward(self, x, want_stats: bool = False, want_tap: bool = False):
        """``want_tap=True`` -> (y, stats_or_None, x_alias), see :func:`ops.depthwise.depthwise_conv2d`."""
        from . import depthwise as DW

        return DW.depthwise_conv2d(x, self.weight, self.bias, self.stride, self.rate, self.padding, self.relu,
                                   self.relu_in, want_stats, want_tap)

    def tf_variables(self):
        d = {f"{self.scope}/depthwise_weights": (self.weight, "depthwise")}
        if self.bias is not None:
            d[f"{self.scope}/biases"] = (self.bias, "plain")
        return d


class SplitSeparableConv2d(TFModule):
    """reference ``split_separable_conv2d`` (core/layers.py:6-49): depthwise 3x3 (bias + ReLU, no BN,
    trunc-normal 0.33) then pointwise 1x1 conv + BN + ReLU (trunc-normal 0.06)."""

    def __init__(self, cin: int, filters: int, kernel_size: int = 3, rate: int = 1, scope: str = "sep",
                 bn: Optional[BNConfig] = None, depthwise_std: float = 0.33, pointwise_std: float = 0.06):
        super().__init__()
        self.depthwise = DepthwiseConv2d(cin, kernel_size, 1, rate, "SAME", True, True, scope=f"{scope}_depthwise",
                                         init_std=depthwise_std)
        self.pointwise = ConvBNAct(cin, filters, 1, scope=f"{scope}_pointwise", bn=bn, init_std=pointwise_std)

    def forward(self, x):
        return self.pointwise(self.depthwise(x))


def max_pool_same(x: torch.Tensor, k: int = 3, stride: int = 2, want_stats: bool = False):
    from . import pooling as P

    return P.max_pool_same(x, k, stride, want_stats)


def global_avg_pool(x: torch.Tensor, keepdims: bool = True) -> torch.Tensor:
    from . import pooling as P

    return P.global_avg_pool(x, keepdims)


def upsample(x: torch.Tensor, out_hw: Tuple[int, int]) -> torch.Tensor:
    from . import resize as RS

    return RS.upsample(x, out_hw)


def fixed_padding(x: torch.Tensor, kernel_size: int, rate: int = 1, mode: str = "constant") -> torch.Tensor:
    return R.fixed_padding(x, kernel_size, rate, mode)


def subsample(x: torch.Tensor, factor: int) -> torch.Tensor:
    return R.subsample(x, factor)
