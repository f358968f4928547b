"""VGG (Simonyan & Zisserman 2014), configuration D = VGG-16 -- BASELINE.json configs[2].
Not present in the reference; written in its style."""
from __future__ import annotations

import torch.nn as nn
from torch import Tensor

from ..nn.layers import AdaptiveAvgPool2d, Conv2d, Dropout, Flatten, Linear, MaxPool2d, fuse_relu_conv, fuse_relu_pool
from ._stem import prepare_input

_CFG_D = (64, 64, "M", 128, 128, "M", 256, 256, 256, "M", 512, 512, 512, "M", 512, 512, 512, "M")


class VGG(nn.Module):
    def __init__(self, cfg=_CFG_D, *, n_classes: int = 1000, p_dropout: float = 0.5, hidden_size: int = 4096,
                 input_mean=None, input_std=None) -> None:
        super().__init__()
        feats = []
        ch = 3
        for v in cfg:
            if v == "M":
                feats.append(MaxPool2d(2, 2))
                if len(feats) > 1 and isinstance(feats[-2], Conv2d):
                    fuse_relu_pool(feats[-2], feats[-1])   # ReLU backward on the pooled-size gradient
            else:
                feats.append(Conv2d(ch, v, 3, 1, 1, relu=True))
                if len(feats) > 1 and isinstance(feats[-2], Conv2d):
                    fuse_relu_conv(feats[-2], feats[-1])   # ReLU backward in the next layer's dgrad epilogue
                ch = v
        self._features = nn.ModuleList(feats)
        if input_mean is not None:   # per-channel image normalisation, fused into the input layout kernel
            self._features[0].set_input_normalization(input_mean, input_std)
        self._avgpool = AdaptiveAvgPool2d((7, 7))
        self._flatten = Flatten()
        self._fc1 = Linear(ch * 7 * 7, hidden_size, relu=True)
        self._drop1 = Dropout(p_dropout)
        self._fc2 = Linear(hidden_size, hidden_size, relu=True)
        self._drop2 = Dropout(p_dropout)
        self._fc3 = Linear(hidden_size, n_classes)

    @property
    def n_classes(self) -> int:
        return self._fc3.out_features

    def forward(self, x: Tensor, /) -> Tensor:
        first = self._features[0]
        x = prepare_input(x, first.weight.dtype)
        for layer in self._features:
            x = layer(x)
        x = self._flatten(self._avgpool(x))
        x = self._drop1(self._fc1(x))
        x = self._drop2(self._fc2(x))
        return self._fc3(x)


class VGG16(VGG):
    def __init__(self, *, n_classes: int = 1000, p_dropout: float = 0.5) -> None:
        super().__init__(_CFG_D, n_classes=n_classes, p_dropout=p_dropout)
