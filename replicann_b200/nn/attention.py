"""
Scaled dot-product self-attention and cross-attention heads and their multihead variants after
https://arxiv.org/abs/1706.03762 -- drop-in for `replicann.nn.attention` (same class names, constructor and
forward signatures, properties, parameter names), running on the library's CUDA kernels: one packed
tcgen05 GEMM for all per-head projections, one fused attention kernel for all heads (no `torch.cat`, no
materialised [B,T,T] scores), one GEMM for the output projection.

Behaviour reproduced from the reference (file:line in /root/reference/src/replicann/nn/attention.py):
  * scale is 1/sqrt(in_dim) -- the embedding size, not the head size (:24);
  * bool mask: True = masked out (-inf) (:39-40); float32 mask: additive (:41-42); anything else -> TypeError
    with the same message (:43-44);
  * `return_kv=True` returns the concatenated (z, k, v) WITHOUT the output projection or its dropout (:183-192);
  * the multi-head wrappers do not forward `p_dropout` to their heads (:80-83), so every head keeps an
    attention-weight dropout of 0.1 in training mode unless given explicitly through `**head_kwargs`;
  * MultiheadCrossAttention(return_kv=True) is broken upstream (:164-165 unpacks a [B,T,hs] tensor over the
    batch dimension -> ValueError unless B == 3); the same ValueError is raised here for every B.
"""
from math import sqrt

import torch
import torch.nn as nn
from torch import Tensor

from ..optim import ParamPack
from . import _functions as Fn
from .layers import Dropout, Linear


def _check_mask(mask: Tensor | None) -> Tensor | None:
    if mask is None:
        return None
    if (mdtype := mask.dtype) == torch.bool:
        m = mask
    elif mdtype == torch.float:
        m = mask
    else:
        raise TypeError(f"Unexpected type `{mdtype}` of the `mask`.")
    return m   # broadcast against [B, T, Tk] (and shape errors) happen in AttentionFn, once q and k are known


def _rng_args(x: Tensor):
    Dropout._global_calls += 1
    seed = torch.cuda.initial_seed() if x.is_cuda else torch.initial_seed()
    return (seed ^ Dropout._seed_mix) & 0xFFFFFFFFFFFFFFFF, Dropout._global_calls


class _AttentionHead(nn.Module):
    def __init__(
        self,
        in_dim: int,
        out_dim: int,
        *,
        bias: bool = False,
        p_dropout: float = 0.1
    ) -> None:
        super().__init__()
        self._scale_coeff = 1 / sqrt(in_dim)
        self._query = Linear(in_dim, out_dim, bias=bias)
        self._dropout = Dropout(p_dropout)

    def _attention(
        self,
        q: Tensor,
        k: Tensor,
        v: Tensor,
        mask: Tensor | None = None,
        *,
        return_kv: bool = False,
    ) -> Tensor:
        mask = _check_mask(mask)
        p = self._dropout.p if self.training else 0.0
        seed, offset = _rng_args(q)
        q3 = q if q.dim() == 3 else q.unsqueeze(0)
        k3 = k if k.dim() == 3 else k.unsqueeze(0)
        v3 = v if v.dim() == 3 else v.unsqueeze(0)
        z = Fn.AttentionFn.apply(q3, k3, v3, mask, self._scale_coeff, p, 1, q3.shape[-1], seed, offset, False)
        z = z if q.dim() == 3 else z.squeeze(0)
        if return_kv:
            return z, k, v
        return z

    @property
    def embeddings_size(self) -> int:
        return self.query.in_features

    @property
    def head_size(self) -> int:
        return self.query.out_features

    @property
    def query(self) -> Linear:
        return self._query


class _MultiheadAttention(nn.Module):
    AttentionHead: type

    def __init__(
        self,
        n_heads: int,
        head_size: int,
        embedding_size: int | None = None,
        *,
        head_bias: bool = False,
        proj_bias: bool = True,
        p_dropout: float = 0.1,
        **head_kwargs,
    ) -> None:
        super().__init__()
        embedding_size = embedding_size or head_size
        self._heads = nn.ModuleList(
            self.AttentionHead(embedding_size, head_size, bias=head_bias, **head_kwargs)
            for _ in range(n_heads)
        )
        self._proj = Linear(n_heads * head_size, embedding_size, bias=proj_bias)
        self._dropout = Dropout(p_dropout)
        # the kernels see ONE [n*H*hs, E] projection matrix; the per-head Linear parameters stay separate
        # nn.Parameters (state_dict parity with the reference) and are declared as a pack so that the flat
        # optimizer lays them out back to back in this order (see optim.ParamPack)
        names = ("_query", "_key", "_value") if hasattr(self._heads[0], "_key") else ("_query",)
        self._wpack = ParamPack.declare([getattr(h, n).weight for n in names for h in self._heads])
        self._bpack = (ParamPack.declare([getattr(h, n).bias for n in names for h in self._heads])
                       if head_bias else None)

    @property
    def embedding_size(self) -> int:
        return self._proj.out_features

    @property
    def head_size(self) -> int:
        return self._proj.in_features // self.n_heads

    @property
    def n_heads(self) -> int:
        return len(self._heads)

    def _project(self, x: Tensor) -> Tensor:
        """All per-head projections of `x` as one GEMM: [.., n*H*hs] with column blocks q_0..q_{H-1} | k_0.. | v_0..
        (what the reference's per-head Linears + torch.cat produce, nn/attention.py:149-192)."""
        ws = self._wpack.members
        bs = self._bpack.members if self._bpack is not None else []
        pack = None
        vw = self._wpack.views()
        if vw is not None:
            vb = self._bpack.views() if self._bpack is not None else (None, None)
            if vb is not None:
                pack = (vw[0], vw[1], vb[0], vb[1])
        return Fn.PackedLinearFn.apply(x, False, len(ws), pack, *ws, *bs)

    def _head_dropout(self) -> float:
        return self._heads[0]._dropout.p if self.training else 0.0


class CrossAttentionHead(_AttentionHead):
    def forward(
        self,
        x: Tensor,
        /,
        k: Tensor,
        v: Tensor,
        mask: Tensor | None = None,
        *,
        return_kv: bool = False,
    ) -> Tensor:
        q = self.query(x)
        return self._attention(q, k, v, mask=mask, return_kv=return_kv)


class SelfAttentionHead(_AttentionHead):
    def __init__(
        self,
        in_dim: int,
        out_dim: int,
        *,
        bias: bool = False,
        p_dropout: float = 0.1
    ) -> None:
        super().__init__(
            in_dim=in_dim,
            out_dim=out_dim,
            bias=bias,
            p_dropout=p_dropout,
        )
        self._key = Linear(in_dim, out_dim, bias=bias)
        self._value = Linear(in_dim, out_dim, bias=bias)

    def forward(self, x: Tensor, /, mask: Tensor | None = None, return_kv: bool = False) -> Tensor:
        q = self.query(x)
        k = self.key(x)
        v = self.value(x)
        return self._attention(q, k, v, mask=mask, return_kv=return_kv)

    @property
    def key(self) -> Linear:
        return self._key

    @property
    def value(self) -> Linear:
        return self._value


class MultiheadCrossAttention(_MultiheadAttention):
    AttentionHead = CrossAttentionHead

    def forward(self, x: Tensor, /, k: Tensor, v: Tensor, mask: Tensor | None = None, *, return_kv: bool = False) -> Tensor:
        if return_kv:
            # upstream defect, see module docstring: `z, k_, v_ = h(x, kp, vp, mask=mask)` unpacks a tensor
            raise ValueError("not enough values to unpack (expected 3, got 2)" if x.shape[0] < 3 else
                             "too many values to unpack (expected 3)" if x.shape[0] > 3 else
                             "MultiheadCrossAttention(return_kv=True) is ill-defined in the reference for batch size 3")
        mask = _check_mask(mask)
        H, hs = self.n_heads, self.head_size
        split_size = self.embedding_size // H
        if split_size != hs:
            raise RuntimeError(f"head_size ({hs}) must equal embedding_size // n_heads ({split_size})")
        q = self._project(x)
        # k / v are used unprojected: head h reads columns [h*split, (h+1)*split) -- the strided layout the
        # kernel takes; only the first H chunks are used, as `torch.split(...)[:H]` does
        seed, offset = _rng_args(x)
        z = Fn.AttentionFn.apply(q, k[..., :H * split_size], v[..., :H * split_size], mask,
                                 self._heads[0]._scale_coeff, self._head_dropout(), H, hs, seed, offset, False)
        return self._dropout(self._proj(z))


class MultiheadSelfAttention(_MultiheadAttention):
    AttentionHead = SelfAttentionHead

    def forward(self, x: Tensor, /, mask: Tensor | None = None, *, return_kv: bool = False) -> Tensor:
        mask = _check_mask(mask)
        H, hs = self.n_heads, self.head_size
        qkv = self._project(x)
        seed, offset = _rng_args(x)
        z = Fn.AttentionFn.apply(qkv, None, None, mask, self._heads[0]._scale_coeff, self._head_dropout(), H, hs,
                                 seed, offset, True)
        if not return_kv:
            return self._dropout(self._proj(z))
        Hh = H * hs
        return z, qkv[..., Hh:2 * Hh].contiguous(), qkv[..., 2 * Hh:].contiguous()
