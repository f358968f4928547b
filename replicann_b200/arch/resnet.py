"""ResNet (He et al., https://arxiv.org/abs/1512.03385), bottleneck variant with the stride on the 3x3
convolution ("v1.5"), written in replicann's block style (arch/transformer.py:44-84: private sub-modules,
keyword-only options, read-only properties).  The reference contains no CNN; this is
the network BASELINE.json's metric is quoted on.  BatchNorm + ReLU (+ residual add) run as one fused kernel.
"""
from __future__ import annotations

import torch
import torch.nn as nn
from torch import Tensor

from ..nn._functions import BnReluMaxPoolFn, BottleneckFn
from ..nn.layers import AdaptiveAvgPool2d, BatchNorm2d, Conv2d, Linear, MaxPool2d
from ._stem import prepare_input


class _Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, in_channels: int, width: int, *, stride: int = 1) -> None:
        super().__init__()
        out_channels = width * self.expansion
        self._conv1 = Conv2d(in_channels, width, 1, bias=False)
        self._bn1 = BatchNorm2d(width, relu=True)
        self._conv2 = Conv2d(width, width, 3, stride, 1, bias=False)
        self._bn2 = BatchNorm2d(width, relu=True)
        self._conv3 = Conv2d(width, out_channels, 1, bias=False)
        self._bn3 = BatchNorm2d(out_channels, relu=True)  # ReLU applied after the residual add
        if stride != 1 or in_channels != out_channels:
            self._down_conv = Conv2d(in_channels, out_channels, 1, stride, bias=False)
            self._down_bn = BatchNorm2d(out_channels)
        else:
            self._down_conv = None
            self._down_bn = None
        self._fused = True

    def forward(self, x: Tensor, /) -> Tensor:
        if self._fused and x.dtype == torch.bfloat16 and x.is_cuda:
            return self._forward_fused(x)
        identity = x if self._down_conv is None else self._down_bn(self._down_conv(x))
        y = self._bn1(self._conv1(x))
        y = self._bn2(self._conv2(y))
        return self._bn3(self._conv3(y), identity)

    def _forward_fused(self, x: Tensor) -> Tensor:
        """Whole block as one autograd node (see BottleneckFn): same kernels, no elementwise gradient adds."""
        bns = [self._bn1, self._bn2, self._bn3] + ([self._down_bn] if self._down_bn is not None else [])
        if self.training:
            for b in bns:
                b._batches_pending += 1
        cfg = tuple((b.eps, b.momentum) for b in bns)
        if len(cfg) == 3:
            cfg = cfg + ((1e-5, 0.1),)
        stats = [t for b in bns for t in (b.running_mean, b.running_var)]
        d = self._down_conv
        return BottleneckFn.apply(x, self._conv2.stride[0], self.training, cfg,
                                  self._conv1.weight, self._bn1.weight, self._bn1.bias,
                                  self._conv2.weight, self._bn2.weight, self._bn2.bias,
                                  self._conv3.weight, self._bn3.weight, self._bn3.bias,
                                  d.weight if d is not None else None,
                                  self._down_bn.weight if d is not None else None,
                                  self._down_bn.bias if d is not None else None, *stats)


class _BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, in_channels: int, width: int, *, stride: int = 1) -> None:
        super().__init__()
        self._conv1 = Conv2d(in_channels, width, 3, stride, 1, bias=False)
        self._bn1 = BatchNorm2d(width, relu=True)
        self._conv2 = Conv2d(width, width, 3, 1, 1, bias=False)
        self._bn2 = BatchNorm2d(width, relu=True)
        if stride != 1 or in_channels != width:
            self._down_conv = Conv2d(in_channels, width, 1, stride, bias=False)
            self._down_bn = BatchNorm2d(width)
        else:
            self._down_conv = None
            self._down_bn = None

    def forward(self, x: Tensor, /) -> Tensor:
        identity = x if self._down_conv is None else self._down_bn(self._down_conv(x))
        y = self._bn1(self._conv1(x))
        return self._bn2(self._conv2(y), identity)


class ResNet(nn.Module):
    def __init__(self, layers: tuple[int, ...] = (3, 4, 6, 3), *, bottleneck: bool = True, n_classes: int = 1000,
                 in_channels: int = 3, base_width: int = 64, input_mean=None, input_std=None) -> None:
        super().__init__()
        block = _Bottleneck if bottleneck else _BasicBlock
        self._conv1 = Conv2d(in_channels, base_width, 7, 2, 3, bias=False)
        if input_mean is not None:   # per-channel image normalisation, fused into the input layout kernel
            self._conv1.set_input_normalization(input_mean, input_std)
        self._bn1 = BatchNorm2d(base_width, relu=True)
        self._maxpool = MaxPool2d(3, 2, 1)
        stages = []
        ch = base_width
        for i, n in enumerate(layers):
            width = base_width * (2 ** i)
            blocks = []
            for j in range(n):
                blocks.append(block(ch, width, stride=(2 if (i > 0 and j == 0) else 1)))
                ch = width * block.expansion
            stages.append(nn.Sequential(*blocks))
        self._stages = nn.ModuleList(stages)
        self._avgpool = AdaptiveAvgPool2d((1, 1))
        self._fc = Linear(ch, n_classes)
        self._n_features = ch
        self._fused_stem = True
        for m in self.modules():
            if isinstance(m, Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")

    @property
    def n_classes(self) -> int:
        return self._fc.out_features

    @property
    def n_features(self) -> int:
        return self._n_features

    def _stem_tail(self, y: Tensor) -> Tensor:
        """BatchNorm + ReLU + max-pool after the first convolution; one pooling kernel applies the normalisation on
        the fly for bf16 CUDA tensors (same values and gradients as the two modules, see BnReluMaxPoolFn)."""
        bn, mp = self._bn1, self._maxpool
        if self._fused_stem and y.is_cuda and y.dtype == torch.bfloat16 and not mp._return_indices:
            if self.training:
                bn._batches_pending += 1
            return BnReluMaxPoolFn.apply(y, bn.weight, bn.bias, bn.running_mean, bn.running_var, self.training,
                                         bn.momentum, bn.eps, mp._k, mp._s, mp._p)
        return mp(bn(y))

    def forward(self, x: Tensor, /) -> Tensor:
        x = prepare_input(x, self._conv1.weight.dtype)
        x = self._stem_tail(self._conv1(x))
        for stage in self._stages:
            x = stage(x)
        x = self._avgpool(x)
        return self._fc(x.reshape(x.shape[0], -1))


class ResNet50(ResNet):
    def __init__(self, *, n_classes: int = 1000) -> None:
        super().__init__((3, 4, 6, 3), bottleneck=True, n_classes=n_classes)


class ResNet18(ResNet):
    def __init__(self, *, n_classes: int = 1000) -> None:
        super().__init__((2, 2, 2, 2), bottleneck=False, n_classes=n_classes)
