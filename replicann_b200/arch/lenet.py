"""LeNet-5 (LeCun et al. 1998), the smallest network BASELINE.json names (configs[0]: fp32 train step,
batch 64, 1x28x28).  Not present in the reference; written in its style.  fp32
weights -> the fp32 CUDA-core kernels (channel counts 1/6/16 are far below a tensor-core tile)."""
from __future__ import annotations

import torch.nn as nn
from torch import Tensor

from ..nn.layers import AvgPool2d, Conv2d, Flatten, Linear, Tanh


class LeNet5(nn.Module):
    def __init__(self, *, n_classes: int = 10, in_channels: int = 1) -> None:
        super().__init__()
        self._conv1 = Conv2d(in_channels, 6, 5, 1, 2)
        self._pool1 = AvgPool2d(2, 2)
        self._conv2 = Conv2d(6, 16, 5)
        self._pool2 = AvgPool2d(2, 2)
        self._flatten = Flatten()
        self._fc1 = Linear(16 * 5 * 5, 120)
        self._fc2 = Linear(120, 84)
        self._fc3 = Linear(84, n_classes)
        self._act = Tanh()

    @property
    def n_classes(self) -> int:
        return self._fc3.out_features

    def forward(self, x: Tensor, /) -> Tensor:
        x = self._pool1(self._act(self._conv1(x)))
        x = self._pool2(self._act(self._conv2(x)))
        x = self._flatten(x)
        x = self._act(self._fc1(x))
        x = self._act(self._fc2(x))
        return self._fc3(x)
