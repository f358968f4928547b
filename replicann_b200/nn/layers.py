"""CNN layer surface named by BASELINE.json's north_star -- Conv2d, Linear, BatchNorm2d, ReLU, MaxPool /
AvgPool, Dropout, softmax-cross-entropy -- as torch.nn.Module subclasses whose constructor arguments,
parameter names / shapes and train()/eval() behaviour match torch.nn (the library the reference composes,
`nn/attention.py:25-26`, `arch/transformer.py:24-26,65`), so `state_dict()`s interchange with torch.nn
modules.  Every forward/backward runs hand-written CUDA through `._functions`; there is no CPU path.

Written in the reference's style (arch/transformer.py:13-41): private `_`-prefixed state, keyword-only
options, read-only properties.
"""
from __future__ import annotations

import math
import os

import torch
import torch.nn as nn
from torch import Tensor

from . import _functions as Fn


def _pair(v):
    return (v, v) if isinstance(v, int) else tuple(v)


class Conv2d(nn.Module):
    """torch.nn.Conv2d(in, out, k, stride, padding, bias) -- groups=1, dilation=1, zeros padding.

    bf16 weights run the tcgen05 implicit-GEMM kernels (input channels % 64 == 0; a first layer with
    fewer channels takes a zero-padded input, see arch/_stem.py); fp32 weights run the fp32 direct kernels.
    """

    def __init__(self, in_channels: int, out_channels: int, kernel_size, stride=1, padding=0, *,
                 bias: bool = True, relu: bool = False, device=None, dtype=None) -> None:
        super().__init__()
        kh, kw = _pair(kernel_size)
        sh, sw = _pair(stride)
        ph, pw = _pair(padding)
        self._stride, self._padding, self._kernel, self._relu = (sh, sw), (ph, pw), (kh, kw), relu
        self._grad_premasked = False     # see fuse_relu_pool()
        self._emit_relu_mask = False     # see fuse_relu_conv()
        w = torch.empty((out_channels, in_channels, kh, kw), device=device, dtype=dtype)
        w = w.contiguous(memory_format=torch.channels_last)
        self.weight = nn.Parameter(w)
        self.bias = nn.Parameter(torch.empty(out_channels, device=device, dtype=dtype)) if bias else None
        self.register_buffer("_in_mean", None, persistent=False)
        self.register_buffer("_in_inv_std", None, persistent=False)
        self.reset_parameters()

    def set_input_normalization(self, mean, std) -> None:
        """For a network's FIRST layer fed with fp32 NCHW images: fold `(x - mean[c]) / std[c]` (torchvision's
        Normalize) into the layout / cast kernel that already touches every input element.  Kept in non-persistent
        buffers, so state_dicts still interchange with torch.nn.Conv2d.  `None, None` removes it."""
        if mean is None:
            self._in_mean = self._in_inv_std = None
            return
        m = torch.as_tensor(mean, dtype=torch.float32).flatten().cpu()
        sd = torch.as_tensor(std, dtype=torch.float32).flatten().cpu()
        if m.numel() != self.in_channels or sd.numel() != self.in_channels:
            raise ValueError(f"mean / std must have {self.in_channels} entries")
        self._in_mean = m.contiguous().to(self.weight.device)
        self._in_inv_std = (1.0 / sd).contiguous().to(self.weight.device)

    def reset_parameters(self) -> None:
        nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))
        if self.bias is not None:
            fan_in = self.weight.shape[1] * self.weight.shape[2] * self.weight.shape[3]
            bound = 1 / math.sqrt(fan_in) if fan_in > 0 else 0
            nn.init.uniform_(self.bias, -bound, bound)

    def _apply(self, fn, recurse=True):
        # keep the filter KRSC (channels_last) through .to()/.cuda() so kernels can read it in place
        super()._apply(fn, recurse)
        w = self.weight
        if not w.permute(0, 2, 3, 1).is_contiguous():
            w.data = w.data.contiguous(memory_format=torch.channels_last)
        if self._in_mean is not None and self._in_mean.dtype != torch.float32:   # the kernels read fp32 constants
            self._in_mean = self._in_mean.float()
            self._in_inv_std = self._in_inv_std.float()
        return self

    @property
    def in_channels(self) -> int:
        return self.weight.shape[1]

    @property
    def out_channels(self) -> int:
        return self.weight.shape[0]

    @property
    def kernel_size(self):
        return self._kernel

    @property
    def stride(self):
        return self._stride

    @property
    def padding(self):
        return self._padding

    def forward(self, x: Tensor, /) -> Tensor:
        w = self.weight
        if (w.dtype == torch.bfloat16 and x.dtype == torch.float32 and self.in_channels <= 8
                and x.shape[1] == self.in_channels and not x.requires_grad):
            # network input: fp32 NCHW images straight into the packed-pixel first-layer kernels
            norm = (self._in_mean, self._in_inv_std) if self._in_mean is not None else None
            emit = self._emit_relu_mask and self._relu and self.out_channels % 64 == 0
            out = Fn.FirstConv2dFn.apply(x, w, self.bias, self._stride, self._padding, self._relu, norm,
                                         self._grad_premasked, emit)
            if emit:
                out[0]._relu_mask = out[1]
                return out[0]
            return out
        if self._in_mean is not None:
            raise RuntimeError("input normalisation is fused into the fp32-image first-layer path only; "
                               "this input does not take it (normalise it upstream)")
        if x.shape[1] != self.in_channels:
            # a channel-padded input (see arch/_stem.prepare_input): zero-pad the filter to match
            if w.dtype == torch.bfloat16 and x.shape[1] > self.in_channels:
                w = Fn.PadFilterFn.apply(w, x.shape[1])
            else:
                raise ValueError(f"expected {self.in_channels} input channels, got {x.shape[1]}")
        in_mask = getattr(x, "_relu_mask", None)      # attached by the producing layer, see fuse_relu_conv()
        if in_mask is not None and (self._stride != (1, 1) or w.dtype != torch.bfloat16 or w is not self.weight):
            raise RuntimeError("fuse_relu_conv: the consumer must be a unit-stride bf16 convolution")
        emit = self._emit_relu_mask and w.dtype == torch.bfloat16 and self.out_channels % 64 == 0
        out = Fn.Conv2dFn.apply(x, w, self.bias, self._stride, self._padding, self._relu, self._grad_premasked,
                                in_mask, emit)
        if emit:
            y, mask = out
            y._relu_mask = mask
            return y
        return out

    def extra_repr(self) -> str:
        return (f"{self.in_channels}, {self.out_channels}, kernel_size={self._kernel}, stride={self._stride}, "
                f"padding={self._padding}, bias={self.bias is not None}")


class Linear(nn.Module):
    """torch.nn.Linear (the one torch layer the reference itself instantiates, nn/attention.py:25,85 and
    arch/transformer.py:24-25).  `relu=True` fuses F.relu (arch/transformer.py:30) into the GEMM epilogue."""

    def __init__(self, in_features: int, out_features: int, *, bias: bool = True, relu: bool = False,
                 device=None, dtype=None) -> None:
        super().__init__()
        self.weight = nn.Parameter(torch.empty((out_features, in_features), device=device, dtype=dtype))
        self.bias = nn.Parameter(torch.empty(out_features, device=device, dtype=dtype)) if bias else None
        self._relu = relu
        self.reset_parameters()

    def reset_parameters(self) -> None:
        nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))
        if self.bias is not None:
            bound = 1 / math.sqrt(self.weight.shape[1])
            nn.init.uniform_(self.bias, -bound, bound)

    @property
    def in_features(self) -> int:
        return self.weight.shape[1]

    @property
    def out_features(self) -> int:
        return self.weight.shape[0]

    def forward(self, x: Tensor, /) -> Tensor:
        return Fn.LinearFn.apply(x, self.weight, self.bias, self._relu)

    def extra_repr(self) -> str:
        return f"{self.in_features}, {self.out_features}, bias={self.bias is not None}, relu={self._relu}"


class BatchNorm2d(nn.Module):
    """torch.nn.BatchNorm2d(num_features, eps, momentum, affine=True, track_running_stats=True).

    Parameters and running statistics stay fp32 whatever dtype the model is cast to (statistics are
    accumulated in fp32 by the kernels).  `relu=True` fuses the following ReLU; `forward(x, residual)`
    fuses the residual add of a ResNet block: y = relu(BN(x) + residual).
    """

    def __init__(self, num_features: int, eps: float = 1e-5, momentum: float = 0.1, *, relu: bool = False,
                 device=None) -> None:
        super().__init__()
        self._eps, self._momentum, self._relu = eps, momentum, relu
        self.weight = nn.Parameter(torch.ones(num_features, device=device, dtype=torch.float32))
        self.bias = nn.Parameter(torch.zeros(num_features, device=device, dtype=torch.float32))
        self.register_buffer("running_mean", torch.zeros(num_features, device=device, dtype=torch.float32))
        self.register_buffer("running_var", torch.ones(num_features, device=device, dtype=torch.float32))
        self.register_buffer("num_batches_tracked", torch.tensor(0, dtype=torch.long, device=device))
        self._batches_pending = 0  # counted on the host; folded into the buffer when the state is read

    def _apply(self, fn, recurse=True):
        super()._apply(fn, recurse)
        for t in (self.weight, self.bias):
            if t.dtype != torch.float32:
                t.data = t.data.float()
        for name in ("running_mean", "running_var"):
            b = getattr(self, name)
            if b.dtype != torch.float32:
                setattr(self, name, b.float())
        return self

    @property
    def num_features(self) -> int:
        return self.weight.shape[0]

    @property
    def eps(self) -> float:
        return self._eps

    @property
    def momentum(self) -> float:
        return self._momentum

    def _flush_counter(self) -> None:
        if self._batches_pending:
            self.num_batches_tracked += self._batches_pending
            self._batches_pending = 0

    def state_dict(self, *args, **kwargs):
        self._flush_counter()
        return super().state_dict(*args, **kwargs)

    def _save_to_state_dict(self, destination, prefix, keep_vars):
        self._flush_counter()
        super()._save_to_state_dict(destination, prefix, keep_vars)

    def forward(self, x: Tensor, /, residual: Tensor | None = None) -> Tensor:
        if self.training:
            self._batches_pending += 1  # no per-step device kernel for a counter nothing on the GPU reads
        return Fn.BatchNormFn.apply(x, self.weight, self.bias, residual, self.running_mean, self.running_var,
                                    self.training, self._momentum, self._eps, self._relu)

    def extra_repr(self) -> str:
        return f"{self.num_features}, eps={self._eps}, momentum={self._momentum}, relu={self._relu}"


class LayerNorm(nn.Module):
    """torch.nn.LayerNorm(E) as used at arch/transformer.py:65,72,188; `forward(x, residual)` computes
    LN(x + residual) in one kernel (the post-LN pattern of arch/transformer.py:111-112)."""

    def __init__(self, normalized_shape: int, eps: float = 1e-5, *, device=None, dtype=None) -> None:
        super().__init__()
        self._eps = eps
        self.weight = nn.Parameter(torch.ones(normalized_shape, device=device, dtype=dtype))
        self.bias = nn.Parameter(torch.zeros(normalized_shape, device=device, dtype=dtype))

    @property
    def normalized_shape(self):
        return (self.weight.shape[0],)

    @property
    def eps(self) -> float:
        return self._eps

    def forward(self, x: Tensor, /, residual: Tensor | None = None) -> Tensor:
        return Fn.LayerNormFn.apply(x, residual, self.weight, self.bias, self._eps)


class Embedding(nn.Module):
    """torch.nn.Embedding(num_embeddings, embedding_dim): weight [V, E], N(0, 1) init, int64 indices.
    `forward(ids, positions=other_embedding)` adds a learned positional table in the same kernel."""

    def __init__(self, num_embeddings: int, embedding_dim: int, *, device=None, dtype=None) -> None:
        super().__init__()
        self.weight = nn.Parameter(torch.empty((num_embeddings, embedding_dim), device=device, dtype=dtype))
        nn.init.normal_(self.weight)

    @property
    def num_embeddings(self) -> int:
        return self.weight.shape[0]

    @property
    def embedding_dim(self) -> int:
        return self.weight.shape[1]

    def forward(self, ids: Tensor, /, positions: "Embedding | None" = None) -> Tensor:
        squeeze = ids.dim() == 1
        ids2 = ids.unsqueeze(0) if squeeze else ids
        if ids2.dim() != 2:
            raise ValueError(f"indices must be [T] or [B, T], got {tuple(ids.shape)}")
        if positions is not None and ids2.shape[1] > positions.num_embeddings:
            raise IndexError(f"sequence length {ids2.shape[1]} exceeds the positional table ({positions.num_embeddings})")
        out = Fn.EmbeddingFn.apply(ids2, self.weight, positions.weight if positions is not None else None)
        return out.squeeze(0) if squeeze else out

    def extra_repr(self) -> str:
        return f"{self.num_embeddings}, {self.embedding_dim}"


class ReLU(nn.Module):
    def __init__(self, inplace: bool = False) -> None:  # `inplace` accepted for signature parity; ignored
        super().__init__()

    def forward(self, x: Tensor, /) -> Tensor:
        return Fn.ReLUFn.apply(x)


class Tanh(nn.Module):
    def forward(self, x: Tensor, /) -> Tensor:
        return Fn.TanhFn.apply(x)


class MaxPool2d(nn.Module):
    """torch.nn.MaxPool2d(kernel_size, stride=None, padding=0, return_indices=False)."""

    def __init__(self, kernel_size, stride=None, padding=0, *, return_indices: bool = False) -> None:
        super().__init__()
        self._k = _pair(kernel_size)
        self._s = _pair(stride) if stride is not None else self._k
        self._p = _pair(padding)
        self._return_indices = return_indices
        self._relu_input = False         # see fuse_relu_pool()

    def forward(self, x: Tensor, /):
        y, idx = Fn.MaxPool2dFn.apply(x, self._k, self._s, self._p, self._return_indices, self._relu_input)
        if self._return_indices:
            return y, idx.long()
        return y


# RNN_FUSE_RELU=0: fuse_relu_pool / fuse_relu_conv declarations are ignored (A/B timing, bit-identity tests)
_FUSE_RELU = os.environ.get("RNN_FUSE_RELU", "1") != "0"


def fuse_relu_pool(conv: "Conv2d", pool: MaxPool2d) -> None:
    """Declare that `pool` is the ONLY consumer of `conv`'s (ReLU-fused) output.  The ReLU backward then runs inside
    the pooling layer's backward on the pooled-size gradient (a window's gradient lands on its argmax element, whose
    value is the pooled value, so the mask `conv_out > 0` at that element equals `pool_out > 0`), and the convolution's
    backward skips its full-size mask pass: its bias gradient becomes a plain column sum (one read of dy instead of
    two reads and a write).  Results are bit-identical to the unfused pair."""
    if not isinstance(conv, Conv2d) or not isinstance(pool, MaxPool2d):
        raise TypeError("fuse_relu_pool(Conv2d, MaxPool2d)")
    if not conv._relu:
        raise ValueError("the convolution has no fused ReLU")
    if _FUSE_RELU:
        conv._grad_premasked = True
        pool._relu_input = True


def fuse_relu_conv(conv: "Conv2d", nxt: "Conv2d") -> None:
    """Declare that `nxt` (unit stride) is the ONLY consumer of `conv`'s (ReLU-fused) output and receives it directly.
    `conv`'s forward then also writes the 1-bit mask (y > 0) from its epilogue and hands it over on the output tensor;
    `nxt` applies it to its input gradient in the epilogue of its dgrad kernel, and `conv`'s backward skips the
    full-size mask pass (its bias gradient becomes a plain column sum).  Bit-identical to the unfused pair.  Takes
    effect for bf16 weights with out_channels % 64 == 0; otherwise the layers run unfused."""
    if not isinstance(conv, Conv2d) or not isinstance(nxt, Conv2d):
        raise TypeError("fuse_relu_conv(Conv2d, Conv2d)")
    if not conv._relu:
        raise ValueError("the convolution has no fused ReLU")
    if nxt._stride != (1, 1):
        raise ValueError("the consumer must have unit stride")
    if _FUSE_RELU:
        conv._emit_relu_mask = True


class AvgPool2d(nn.Module):
    """torch.nn.AvgPool2d(kernel_size, stride=None, padding=0) with torch's defaults otherwise."""

    def __init__(self, kernel_size, stride=None, padding=0) -> None:
        super().__init__()
        self._k = _pair(kernel_size)
        self._s = _pair(stride) if stride is not None else self._k
        self._p = _pair(padding)

    def forward(self, x: Tensor, /) -> Tensor:
        return Fn.AvgPool2dFn.apply(x, self._k, self._s, self._p)


class AdaptiveAvgPool2d(nn.Module):
    """torch.nn.AdaptiveAvgPool2d for the output sizes the classic nets use: (1, 1), or an output size
    that divides the input evenly (then it is an AvgPool2d with kernel = stride = in/out)."""

    def __init__(self, output_size) -> None:
        super().__init__()
        self._out = _pair(output_size)

    def forward(self, x: Tensor, /) -> Tensor:
        H, W = x.shape[2], x.shape[3]
        oh, ow = self._out
        if (oh, ow) == (1, 1):
            return Fn.GlobalAvgPoolFn.apply(x)
        if (oh, ow) == (H, W):
            return x
        if H % oh or W % ow:
            raise ValueError(f"AdaptiveAvgPool2d: input {H}x{W} is not a multiple of output {oh}x{ow}")
        k = (H // oh, W // ow)
        return Fn.AvgPool2dFn.apply(x, k, k, (0, 0))


class Dropout(nn.Module):
    """torch.nn.Dropout(p).  Counter-based Philox stream keyed by (torch's CUDA seed, per-module call
    counter): reproducible under torch.manual_seed, but a different stream from ATen's (parity tests run
    with p = 0 or .eval(), as SURVEY.md section 4 prescribes)."""

    _global_calls = 0
    _seed_mix = 0   # set per rank by parallel.DataParallel so that replicas draw different masks

    def __init__(self, p: float = 0.5, inplace: bool = False) -> None:
        super().__init__()
        if not 0.0 <= p < 1.0 and p != 1.0:
            raise ValueError(f"dropout probability has to be between 0 and 1, but got {p}")
        self._p = float(p)

    @property
    def p(self) -> float:
        return self._p

    def forward(self, x: Tensor, /) -> Tensor:
        if not self.training or self._p == 0.0:
            return x
        Dropout._global_calls += 1
        seed = torch.cuda.initial_seed() if x.is_cuda else torch.initial_seed()
        return Fn.DropoutFn.apply(x, self._p, (seed ^ Dropout._seed_mix) & 0xFFFFFFFFFFFFFFFF, Dropout._global_calls)

    def extra_repr(self) -> str:
        return f"p={self._p}"


class Flatten(nn.Module):
    """torch.nn.Flatten(1): NCHW-ordered flatten, so nn.Linear weights keep torch's column order."""

    def forward(self, x: Tensor, /) -> Tensor:
        if x.dim() == 4 and x.shape[2] * x.shape[3] > 1:
            return x.contiguous().reshape(x.shape[0], -1)  # layout change (data movement only)
        return x.reshape(x.shape[0], -1)


class CrossEntropyLoss(nn.Module):
    """torch.nn.CrossEntropyLoss() (mean reduction) -- fused softmax + NLL + gradient kernel."""

    def forward(self, logits: Tensor, target: Tensor) -> Tensor:
        return Fn.SoftmaxCrossEntropyFn.apply(logits, target)


def cross_entropy(logits: Tensor, target: Tensor) -> Tensor:
    return Fn.SoftmaxCrossEntropyFn.apply(logits, target)
