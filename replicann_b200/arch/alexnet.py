"""AlexNet (Krizhevsky et al. 2012, single-tower torchvision layout) -- BASELINE.json configs[1].
Not present in the reference; written in its style.  Convolution bias + ReLU are
fused into the tcgen05 epilogue, Linear bias + ReLU into the GEMM epilogue."""
from __future__ import annotations

import torch.nn as nn
from torch import Tensor

from ..nn.layers import AdaptiveAvgPool2d, Conv2d, Dropout, Flatten, Linear, MaxPool2d, fuse_relu_conv, fuse_relu_pool
from ._stem import prepare_input


class AlexNet(nn.Module):
    def __init__(self, *, n_classes: int = 1000, p_dropout: float = 0.5, input_mean=None, input_std=None) -> None:
        super().__init__()
        self._conv1 = Conv2d(3, 64, 11, 4, 2, relu=True)
        if input_mean is not None:   # per-channel image normalisation, fused into the input layout kernel
            self._conv1.set_input_normalization(input_mean, input_std)
        self._pool1 = MaxPool2d(3, 2)
        self._conv2 = Conv2d(64, 192, 5, 1, 2, relu=True)
        self._pool2 = MaxPool2d(3, 2)
        self._conv3 = Conv2d(192, 384, 3, 1, 1, relu=True)
        self._conv4 = Conv2d(384, 256, 3, 1, 1, relu=True)
        self._conv5 = Conv2d(256, 256, 3, 1, 1, relu=True)
        self._pool5 = MaxPool2d(3, 2)
        for conv, pool in ((self._conv1, self._pool1), (self._conv2, self._pool2), (self._conv5, self._pool5)):
            fuse_relu_pool(conv, pool)                     # ReLU backward on the pooled-size gradient
        fuse_relu_conv(self._conv3, self._conv4)           # ReLU backward in the next layer's dgrad epilogue
        fuse_relu_conv(self._conv4, self._conv5)
        self._avgpool = AdaptiveAvgPool2d((6, 6))
        self._flatten = Flatten()
        self._drop1 = Dropout(p_dropout)
        self._fc1 = Linear(256 * 6 * 6, 4096, relu=True)
        self._drop2 = Dropout(p_dropout)
        self._fc2 = Linear(4096, 4096, relu=True)
        self._fc3 = Linear(4096, n_classes)

    @property
    def n_classes(self) -> int:
        return self._fc3.out_features

    def forward(self, x: Tensor, /) -> Tensor:
        x = prepare_input(x, self._conv1.weight.dtype)
        x = self._pool1(self._conv1(x))
        x = self._pool2(self._conv2(x))
        x = self._conv5(self._conv4(self._conv3(x)))
        x = self._flatten(self._avgpool(self._pool5(x)))
        x = self._fc1(self._drop1(x))
        x = self._fc2(self._drop2(x))
        return self._fc3(x)
