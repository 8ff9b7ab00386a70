"""Conv1d/Linear blocks with the parameter names of the reference's deepclr/models/helper.py
(`_sequential.{i}._sequential.0.{weight,bias}`), xavier-uniform weights, zero bias, ReLU after every layer."""
from typing import List

import torch
from torch import nn
import torch.nn.functional as F


class _Block(nn.Module):
    def __init__(self, layer: nn.Module, norm: nn.Module = None):
        super().__init__()
        nn.init.xavier_uniform_(layer.weight)
        if layer.bias is not None:
            layer.bias.data.fill_(0.0)
        self._sequential = nn.Sequential(layer) if norm is None else nn.Sequential(layer, norm)
        self._output_dim = layer.weight.shape[0]

    def output_dim(self) -> int:
        return self._output_dim

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return F.relu(self._sequential(x))


class Conv1d(_Block):
    def __init__(self, in_channels: int, out_channels: int, kernel_size=1, bias: bool = True,
                 batch_norm: bool = False):
        super().__init__(nn.Conv1d(in_channels, out_channels, kernel_size=kernel_size, bias=bias),
                         nn.BatchNorm1d(out_channels) if batch_norm else None)


class Linear(_Block):
    def __init__(self, in_features: int, out_features: int, bias: bool = True, batch_norm: bool = False):
        super().__init__(nn.Linear(in_features, out_features, bias=bias),
                         nn.BatchNorm1d(out_features) if batch_norm else None)


class _MultiLayer(nn.Module):
    def __init__(self, make, layer_sizes: List[int], batch_norm: bool, dropout_keep: float, dropout_last: bool):
        super().__init__()
        layers: List[nn.Module] = []
        pairs = list(zip(layer_sizes[:-1], layer_sizes[1:]))
        for i, (size_in, size_out) in enumerate(pairs):
            layers.append(make(size_in, size_out, batch_norm))
            last = i == len(pairs) - 1
            if dropout_keep < 1.0 and (not last or dropout_last):
                layers.append(nn.Dropout(1.0 - dropout_keep))
        self._sequential = nn.Sequential(*layers)
        self._output_dim = layer_sizes[-1]

    def output_dim(self) -> int:
        return self._output_dim

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self._sequential(x)


class Conv1dMultiLayer(_MultiLayer):
    def __init__(self, layer_sizes: List[int], batch_norm: bool = False, dropout_keep: float = 1.0,
                 dropout_last: bool = False):
        super().__init__(lambda i, o, bn: Conv1d(i, o, 1, bias=True, batch_norm=bn), layer_sizes, batch_norm,
                         dropout_keep, dropout_last)


class LinearMultiLayer(_MultiLayer):
    def __init__(self, layer_sizes: List[int], batch_norm: bool = False, dropout_keep: float = 1.0,
                 dropout_last: bool = False):
        super().__init__(lambda i, o, bn: Linear(i, o, bias=True, batch_norm=bn), layer_sizes, batch_norm,
                         dropout_keep, dropout_last)
