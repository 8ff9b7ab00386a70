"""SharedMLP with the parameter names of the reference's pytorch_utils.py (:5-32, :163-197):
`layer{i}.conv.{weight,bias}` (+ `layer{i}.bn.bn.*` when bn=True), kaiming-normal weights, zero bias."""
from typing import List

import torch.nn as nn


class _BN2d(nn.Sequential):
    def __init__(self, channels: int):
        super().__init__()
        self.add_module('bn', nn.BatchNorm2d(channels))
        nn.init.constant_(self[0].weight, 1.0)
        nn.init.constant_(self[0].bias, 0)


class Conv2d(nn.Sequential):
    def __init__(self, in_size: int, out_size: int, *, bn: bool = False, activation=nn.ReLU(inplace=True)):
        super().__init__()
        conv = nn.Conv2d(in_size, out_size, kernel_size=(1, 1), bias=not bn)
        nn.init.kaiming_normal_(conv.weight)
        if conv.bias is not None:
            nn.init.constant_(conv.bias, 0)
        self.add_module('conv', conv)
        if bn:
            self.add_module('bn', _BN2d(out_size))
        if activation is not None:
            self.add_module('activation', activation)


class SharedMLP(nn.Sequential):
    def __init__(self, args: List[int], *, bn: bool = False, activation=nn.ReLU(inplace=True)):
        super().__init__()
        for i in range(len(args) - 1):
            self.add_module('layer{}'.format(i), Conv2d(args[i], args[i + 1], bn=bn, activation=activation))
