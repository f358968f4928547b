"""
Transformer blocks after https://arxiv.org/abs/1706.03762 -- drop-in for `replicann.arch.transformer`
(same class names, signatures, properties, state_dict keys incl. the `_attn_mask` buffer), on the library's
CUDA kernels: tcgen05 GEMMs with the FFN's bias+ReLU fused into the epilogue, fused attention, and one
kernel for each `LayerNorm(x + sublayer(x))`.

Reproduced reference behaviour (file:line in /root/reference/src/replicann/arch/transformer.py):
  * post-LN wiring (:111-112, :149-150, :198-200); FFN = Linear, relu, Linear, Dropout (:28-33);
  * `return_kv=True` adds the UNPROJECTED attention output to the residual (:114-115);
  * the decoders' `_attn_mask` is a float32 tril(ones) buffer (:142-145,176-179) and therefore takes the
    ADDITIVE mask branch -- the blocks are not causal.  After `.to(torch.bfloat16)` the buffer
    is bf16 and forward raises the reference's TypeError, exactly as upstream does.
"""
import torch
import torch.nn as nn
from torch import Tensor

from replicann_b200.nn.attention import MultiheadCrossAttention, MultiheadSelfAttention
from replicann_b200.nn.layers import Dropout, LayerNorm, Linear


class _TransformerFFN(nn.Module):
    def __init__(
        self,
        embedding_size: int,
        *,
        bias: bool = True,
        hidden_size: int | None = None,
        p_dropout: float = 0.1,
    ) -> None:
        super().__init__()
        hidden_size = hidden_size or 4 * embedding_size
        self._upscale = Linear(embedding_size, hidden_size, bias=bias, relu=True)  # F.relu fused (:30)
        self._downscale = Linear(hidden_size, embedding_size, bias=bias)
        self._dropout = Dropout(p_dropout)

    def forward(self, x: Tensor, /) -> None:
        x = self._upscale(x)
        x = self._downscale(x)
        x = self._dropout(x)
        return x

    @property
    def embedding_size(self) -> int:
        return self._upscale.in_features

    @property
    def hidden_size(self) -> int:
        return self._upscale.out_features


class _TransformerBlock(nn.Module):
    def __init__(
        self,
        n_heads: int,
        embedding_size: int,
        *,
        ffn_bias: bool = True,
        ffn_hidden_size: int | None = None,
        head_bias: bool = False,
        proj_bias: bool = True,
        p_dropout: float = 0.1,
    ) -> None:
        super().__init__()
        self._attn = MultiheadSelfAttention(
            n_heads,
            embedding_size // n_heads,
            embedding_size,
            head_bias=head_bias,
            proj_bias=proj_bias,
            p_dropout=p_dropout,
        )
        self._attn_ln = LayerNorm(embedding_size)
        self._ffn = _TransformerFFN(
            embedding_size,
            bias=ffn_bias,
            hidden_size=ffn_hidden_size,
            p_dropout=p_dropout,
        )
        self._ffn_ln = LayerNorm(embedding_size)

    @property
    def embedding_size(self) -> int:
        return self._attn.embedding_size

    @property
    def head_size(self) -> int:
        return self._attn.head_size

    @property
    def n_heads(self) -> int:
        return self._attn.n_heads


class TransformerEncoder(_TransformerBlock):
    def __init__(
        self,
        n_heads: int,
        embedding_size: int,
        *,
        ffn_bias: bool = True,
        ffn_hidden_size: int | None = None,
        head_bias: bool = False,
        proj_bias: bool = True,
        p_dropout: float = 0.1,
    ) -> None:
        super().__init__(
            n_heads=n_heads,
            embedding_size=embedding_size,
            ffn_bias=ffn_bias,
            ffn_hidden_size=ffn_hidden_size,
            head_bias=head_bias,
            proj_bias=proj_bias,
            p_dropout=p_dropout,
        )

    def forward(self, x: Tensor, *, return_kv: bool = False) -> Tensor:
        if not return_kv:
            x = self._attn_ln(self._attn(x), x)          # LN(x + attn(x))
            x = self._ffn_ln(self._ffn(x), x)            # LN(x + ffn(x))
            return x
        z, k, v = self._attn(x, return_kv=True)
        x = self._attn_ln(z, x)
        x = self._ffn_ln(self._ffn(x), x)
        return x, k, v


class TransformerDecoder(_TransformerBlock):
    def __init__(
        self,
        n_heads: int,
        embedding_size: int,
        *,
        context_size: int,
        ffn_bias: bool = True,
        ffn_hidden_size: int | None = None,
        head_bias: bool = False,
        proj_bias: bool = True,
        p_dropout: float = 0.1,
    ) -> None:
        super().__init__(
            n_heads=n_heads,
            embedding_size=embedding_size,
            ffn_bias=ffn_bias,
            ffn_hidden_size=ffn_hidden_size,
            head_bias=head_bias,
            proj_bias=proj_bias,
            p_dropout=p_dropout,
        )
        self.register_buffer(
            "_attn_mask",
            torch.tril(torch.ones(context_size, context_size)),
        )

    def forward(self, x: Tensor) -> Tensor:
        n_tokens = x.shape[1]
        x = self._attn_ln(self._attn(x, mask=self._attn_mask[:n_tokens, :n_tokens]), x)
        x = self._ffn_ln(self._ffn(x), x)
        return x


class TransformerCrossDecoder(_TransformerBlock):
    def __init__(
        self,
        n_heads: int,
        embedding_size: int,
        *,
        context_size: int,
        ffn_bias: bool = True,
        ffn_hidden_size: int | None = None,
        head_bias: bool = False,
        proj_bias: bool = True,
        p_dropout: float = 0.1,
    ) -> None:
        super().__init__(
            n_heads=n_heads,
            embedding_size=embedding_size,
            ffn_bias=ffn_bias,
            ffn_hidden_size=ffn_hidden_size,
            head_bias=head_bias,
            proj_bias=proj_bias,
            p_dropout=p_dropout,
        )
        self.register_buffer(
            "_attn_mask",
            torch.tril(torch.ones(context_size, context_size)),
        )
        self._cross_attn = MultiheadCrossAttention(
            n_heads,
            embedding_size // n_heads,
            embedding_size,
            head_bias=head_bias,
            proj_bias=proj_bias,
            p_dropout=p_dropout,
        )
        self._cross_attn_ln = LayerNorm(embedding_size)

    def forward(
        self,
        x: Tensor,
        /,
        encoder_k: Tensor,
        encoder_v: Tensor,
    ) -> Tensor:
        n_tokens = x.shape[1]
        x = self._attn_ln(self._attn(x, mask=self._attn_mask[:n_tokens, :n_tokens]), x)
        x = self._cross_attn_ln(self._cross_attn(x, encoder_k, encoder_v), x)
        x = self._ffn_ln(self._ffn(x), x)
        return x
