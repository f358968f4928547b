"""Input stage shared by the ImageNet-shaped CNNs: fp32 NCHW images -> bf16 NHWC with the channel
dimension zero-padded to one 64-wide tensor-core K step, and the matching zero-padded first filter."""
from __future__ import annotations

import torch
from torch import Tensor

from .. import functional as RF
from ..nn import _functions as Fn

STEM_CPAD = 64


def prepare_input(x: Tensor, weight_dtype, cpad: int = STEM_CPAD) -> Tensor:
    """Bring a user batch to the layout the first convolution reads.

    fp32 NCHW-contiguous images for a bf16 model go through the layout/cast kernel (the input pipeline
    step of SURVEY.md section 8f.2); tensors that already match pass through untouched."""
    if weight_dtype == torch.bfloat16:
        if x.dtype == torch.float32:
            if x.shape[1] <= 8 and not x.requires_grad:
                return x  # the first Conv2d consumes fp32 NCHW images directly (packed-pixel kernels)
            if not x.is_contiguous():
                x = x.contiguous()
            return Fn.nchw(RF.nchw_to_nhwc_bf16(x, cpad=cpad))
        if x.dtype == torch.bfloat16:
            return x
        raise TypeError("bf16 models take fp32 NCHW images (or a bf16 channels_last tensor)")
    if x.dtype != torch.float32:
        raise TypeError(f"fp32 model got input of dtype {x.dtype}")
    return x
