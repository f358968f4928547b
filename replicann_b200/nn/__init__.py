"""Layer surface: the reference's attention layers (`replicann.nn.attention`) plus the CNN layers the
north star names.  Import by full module path like the reference does (its __init__ files are empty);
the names below are re-exported for convenience."""
from .layers import (AdaptiveAvgPool2d, AvgPool2d, BatchNorm2d, Conv2d, CrossEntropyLoss, Dropout, Flatten,
                     LayerNorm, Linear, MaxPool2d, ReLU, Tanh, cross_entropy)

__all__ = ["AdaptiveAvgPool2d", "AvgPool2d", "BatchNorm2d", "Conv2d", "CrossEntropyLoss", "Dropout", "Flatten",
           "LayerNorm", "Linear", "MaxPool2d", "ReLU", "Tanh", "cross_entropy"]
