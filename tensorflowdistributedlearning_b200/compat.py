"""The reference's *functional* call signatures over this package's modules.

The reference is TensorFlow-1 graph code: ``slim`` layer functions create their variables on first use inside a
``tf.variable_scope`` and return tensors; streaming metrics return ``(value, update_op)`` pairs backed by local
variables.  The package itself is built from ``nn.Module`` objects.  This file is the bridge for a user who calls the
reference's functions the reference's way:

* a **variable store** keyed by scope name plays the part of TF's variable scopes: the first call under a scope
  builds the module (shapes inferred from the input), later calls reuse it (``reuse=True`` semantics);
  ``variables(scope)`` / ``reset_variables()`` expose and clear it
* ``data_format="NCHW"`` arguments are honoured at this edge by a permute in and out (the internal layout is NHWC)
* metric functions keep ``total`` / ``count`` accumulators per ``name`` (TF local variables);
  ``reset_local_variables()`` is ``tf.local_variables_initializer()``

reference: core/layers.py:6-109, core/resnet.py:57-152,398-496, core/metric.py:6-71, core/losses.py:18-92.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import torch

from .ops import functional as TF
from .ops import layers as L
from .ops import reference as R

_VARIABLES: Dict[str, torch.nn.Module] = {}
_LOCALS: Dict[str, "_MeanState"] = {}


# ---------------------------------------------------------------------------------------------
# variable store
# ---------------------------------------------------------------------------------------------
def variables(scope: Optional[str] = None):
    """The module created under ``scope`` (or the whole store)."""
    return _VARIABLES if scope is None else _VARIABLES[scope]


def reset_variables() -> None:
    _VARIABLES.clear()


def _get(scope: str, build, like: torch.Tensor, is_training: Optional[bool] = None) -> torch.nn.Module:
    m = _VARIABLES.get(scope)
    if m is None:
        m = build().to(like.device)
        _VARIABLES[scope] = m
    if is_training is not None:
        m.train(bool(is_training))
    return m


def _to_nhwc(x: torch.Tensor, data_format: str) -> torch.Tensor:
    if data_format == "NCHW":
        return x.permute(0, 2, 3, 1).contiguous()
    if data_format != "NHWC":
        raise ValueError(f"Unknown data format {data_format}. Has to be either NCHW or NHWC")
    return x


def _from_nhwc(x: torch.Tensor, data_format: str) -> torch.Tensor:
    return x.permute(0, 3, 1, 2).contiguous() if data_format == "NCHW" else x


# ---------------------------------------------------------------------------------------------
# core/layers.py
# ---------------------------------------------------------------------------------------------
def split_separable_conv2d(inputs: torch.Tensor, filters: int, kernel_size: int = 3, rate: int = 1,
                           weight_decay: float = 0.00004, depthwise_weights_initializer_stddev: float = 0.33,
                           pointwise_weights_initializer_stddev: float = 0.06, scope: Optional[str] = None,
                           data_format: str = "NHWC", is_training: Optional[bool] = None) -> torch.Tensor:
    """core/layers.py:6-49: depthwise kxk (bias + ReLU, no BN) then pointwise 1x1 conv + BN + ReLU; variables live
    under ``{scope}_depthwise`` / ``{scope}_pointwise``."""
    if scope is None:
        raise ValueError("split_separable_conv2d needs a scope (the reference concatenates it into variable names)")
    x = _to_nhwc(inputs, data_format)
    m = _get(scope, lambda: L.SplitSeparableConv2d(
        x.shape[-1], filters, kernel_size, rate, scope=scope, bn=L.BNConfig(weight_decay=weight_decay),
        depthwise_std=depthwise_weights_initializer_stddev, pointwise_std=pointwise_weights_initializer_stddev),
        x, is_training)
    return _from_nhwc(m(x), data_format)


def _fixed_padding(inputs: torch.Tensor, kernel_size: int, data_format: str = "NCHW", mode: str = "CONSTANT"):
    """core/layers.py:52-79 (note the reference's default data_format here is NCHW)."""
    return _from_nhwc(R.fixed_padding(_to_nhwc(inputs, data_format), kernel_size, 1, mode.lower()), data_format)


def _upsample(inputs: torch.Tensor, out_shape: Sequence[int], data_format: str = "NCHW") -> torch.Tensor:
    """core/layers.py:82-109: symmetric pad 1, legacy bilinear resize to (H+4, W+4), crop 2.  The reference reads
    ``height = out_shape[1]``, ``width = out_shape[0]`` (:96-99); so does this."""
    return _from_nhwc(L.upsample(_to_nhwc(inputs, data_format), (int(out_shape[1]), int(out_shape[0]))), data_format)


# ---------------------------------------------------------------------------------------------
# core/resnet.py
# ---------------------------------------------------------------------------------------------
def _unit(kind: str, inputs, depth, depth_bottleneck, stride, unit_rate, rate, outputs_collections, scope,
          data_format, is_training):
    from .models.resnet import UNIT_FNS

    if scope is None:
        raise ValueError("a scope is required")
    x = _to_nhwc(inputs, data_format)
    m = _get(scope, lambda: UNIT_FNS[kind](x.shape[-1], depth, depth_bottleneck, stride, rate * unit_rate, scope,
                                            L.BNConfig()), x, is_training)
    out = m(x)
    if isinstance(outputs_collections, dict):
        outputs_collections[m.scope] = out
    elif isinstance(outputs_collections, list):
        outputs_collections.append(out)
    return _from_nhwc(out, data_format)


def bottleneck(inputs, depth, depth_bottleneck, stride, unit_rate=1, rate=1, outputs_collections=None, scope=None,
               data_format: str = "NHWC", is_training: Optional[bool] = None):
    """core/resnet.py:94-152 (the output is collected under ``{scope}/bottleneck_v2`` like slim does)."""
    return _unit("bottleneck", inputs, depth, depth_bottleneck, stride, unit_rate, rate, outputs_collections, scope,
                 data_format, is_training)


def basic_block(inputs, depth, depth_bottleneck, stride, unit_rate=1, rate=1, outputs_collections=None, scope=None,
                data_format: str = "NHWC", is_training: Optional[bool] = None):
    """core/resnet.py:57-91."""
    return _unit("basic_block", inputs, depth, depth_bottleneck, stride, unit_rate, rate, outputs_collections, scope,
                 data_format, is_training)


def resnet_model(input, model_name, weight_decay, batch_norm_decay, batch_norm_epsilon, batch_norm_scale, data_format,
                 is_training, output_stride, base_depth, input_shape, n_blocks, block_type):  # noqa: A002
    """core/resnet.py:398-496 with the reference's positional signature: returns the logits (pre-sigmoid) in
    ``data_format``.  Variables live under ``model_name`` (model.py:353)."""
    from .models.resnet import ResNetSegModel

    x = _to_nhwc(input, data_format)
    m = _get(model_name, lambda: ResNetSegModel(
        model_name=model_name, in_channels=x.shape[-1], weight_decay=weight_decay, batch_norm_decay=batch_norm_decay,
        batch_norm_epsilon=batch_norm_epsilon, batch_norm_scale=batch_norm_scale, output_stride=output_stride,
        base_depth=base_depth, input_shape=tuple(input_shape), n_blocks=tuple(n_blocks), block_type=block_type),
        x, is_training)
    if x.is_cuda and x.dtype == torch.float32:
        x = x.to(torch.bfloat16)  # the native kernels' activation dtype
    return _from_nhwc(m(x), data_format)


# ---------------------------------------------------------------------------------------------
# core/metric.py
# ---------------------------------------------------------------------------------------------
class _MeanState:
    def __init__(self):
        self.total, self.count = 0.0, 0.0

    def value(self) -> float:
        return self.total / self.count if self.count > 0 else 0.0


def reset_local_variables(name: Optional[str] = None) -> None:
    """``tf.local_variables_initializer()`` for the streaming metrics (all of them, or the one called ``name``)."""
    if name is None:
        _LOCALS.clear()
    else:
        _LOCALS.pop(name, None)


def _streaming_mean(scores: torch.Tensor, name: str, metrics_collections, updates_collections) -> Tuple[float, float]:
    st = _LOCALS.setdefault(name, _MeanState())
    value = st.value()                      # `iou`: the ratio BEFORE this batch is added
    st.total += float(scores.double().sum())
    st.count += scores.numel()
    update = st.value()                     # `update_op`: adds the batch, returns the new ratio (A22)
    if metrics_collections is not None:
        metrics_collections.append(value)
    if updates_collections is not None:
        updates_collections.append(update)
    return value, update


def mIOU(y_true: torch.Tensor, y_pred: torch.Tensor, metrics_collections: Optional[List] = None,
         updates_collections: Optional[List] = None, name: str = "iou") -> Tuple[float, float]:
    """core/metric.py:6-50: ``y_pred`` is the thresholded 0/1 prediction (model.py:372,397).  Returns
    ``(value, update_op)`` of the streaming mean of the per-image thresholded score (with the reference's extra
    ``score *`` factor, defect 14)."""
    B = y_true.shape[0]
    logits = y_pred.reshape(B, -1).float() * 2.0 - 1.0  # > 0 exactly where the prediction is 1
    scores, _ = TF.seg_metrics(logits, y_true.reshape(B, -1).float(), 0.5, True)
    return _streaming_mean(scores, name, metrics_collections, updates_collections)


def mean_accuracy(y_true: torch.Tensor, y_pred: torch.Tensor, metrics_collections: Optional[List] = None,
                  updates_collections: Optional[List] = None, name: str = "acc") -> Tuple[float, float]:
    """core/metric.py:53-71: streaming mean of the per-image fraction of pixels where ``y_true == y_pred``."""
    B = y_true.shape[0]
    scores = (y_true.reshape(B, -1) == y_pred.reshape(B, -1)).float().mean(dim=1)
    return _streaming_mean(scores, name, metrics_collections, updates_collections)


# ---------------------------------------------------------------------------------------------
# core/losses.py
# ---------------------------------------------------------------------------------------------
def flatten_binary_scores(scores: torch.Tensor, labels: torch.Tensor, ignore=None):
    """core/losses.py:68-80: flatten, and drop the positions whose label equals ``ignore``."""
    scores, labels = scores.reshape(-1), labels.reshape(-1)
    if ignore is None:
        return scores, labels
    valid = labels != ignore
    return scores[valid], labels[valid]


def lovasz_hinge(logits: torch.Tensor, labels: torch.Tensor, per_image: bool = True, ignore=None) -> torch.Tensor:
    """core/losses.py:18-37.  Without ``ignore`` the per-image form runs on the GPU kernel for CUDA inputs; with it the
    images have different lengths after masking and are reduced one by one with the torch restatement."""
    if per_image and ignore is None:
        return TF.lovasz_loss(labels, logits)
    if per_image:
        losses = [R.lovasz_hinge_flat(*(t.float() for t in flatten_binary_scores(lg, lb, ignore)))
                  for lg, lb in zip(logits, labels)]
        return torch.stack(losses).mean()
    lg, lb = flatten_binary_scores(logits, labels, ignore)
    return R.lovasz_hinge_flat(lg.float(), lb.float())


def lovasz_loss(y_true: torch.Tensor, y_pred: torch.Tensor, data_format: str = "NHWC") -> torch.Tensor:
    """core/losses.py:83-92: squeeze the channel axis (-1 for NHWC, 1 for NCHW), per-image hinge, batch mean."""
    ax = 1 if data_format == "NCHW" else -1
    return lovasz_hinge(y_pred.squeeze(ax), y_true.squeeze(ax), per_image=True, ignore=None)
