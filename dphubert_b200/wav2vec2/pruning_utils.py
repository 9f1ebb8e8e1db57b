"""In-place structural slicers used by `prune()` — mirror of the reference `wav2vec2/pruning_utils.py:9-51`.

Cold host-side path (the reference prunes once, offline, on CPU); results must be bit-exact in shapes and kept
indices, so these are plain `index_select`s on the master fp32 parameters.
"""
from typing import Union

import torch
import torch.nn as nn


def _axis(dim: str) -> int:
    if dim == "input":
        return 1
    if dim == "output":
        return 0
    raise ValueError(dim)


def _take(p: torch.Tensor, axis: int, index: torch.Tensor, requires_grad: bool = True) -> nn.Parameter:
    return nn.Parameter(p.index_select(axis, index).clone().detach(), requires_grad=requires_grad)


def prune_linear_layer(layer: nn.Linear, index: torch.LongTensor, dim: str):
    """Keep rows (`dim="output"`) or columns (`dim="input"`) `index` of a Linear, in place."""
    axis = _axis(dim)
    if axis == 1:
        layer.in_features = len(index)
    else:
        layer.out_features = len(index)
    layer.weight = _take(layer.weight, axis, index)
    if layer.bias is not None and axis == 0:
        layer.bias = _take(layer.bias, 0, index)


def prune_conv1d_layer(layer: nn.Conv1d, index: torch.LongTensor, dim: str):
    """Keep output / input channels `index` of a Conv1d (weight is (out, in, k)), in place."""
    axis = _axis(dim)
    if axis == 1:
        layer.in_channels = len(index)
    else:
        layer.out_channels = len(index)
    layer.weight = _take(layer.weight, axis, index)
    if layer.bias is not None and axis == 0:
        layer.bias = _take(layer.bias, 0, index)


def prune_layer_norm(layernorm: Union[nn.LayerNorm, nn.GroupNorm], index: torch.LongTensor):
    """Keep affine entries `index` of a LayerNorm / per-channel GroupNorm, in place."""
    layernorm.weight = _take(layernorm.weight, 0, index)
    layernorm.bias = _take(layernorm.bias, 0, index)
    n = len(index)
    if isinstance(layernorm, nn.LayerNorm):
        layernorm.normalized_shape = (n,)
    elif isinstance(layernorm, nn.GroupNorm):
        layernorm.num_groups = n
        layernorm.num_channels = n
