"""
Encoder / decoder stacks, token + positional embeddings, and a language-model head with its loss -- the networks
the reference's blocks are the building bricks of (https://arxiv.org/abs/1706.03762, section 3).

The reference stops at single blocks (/root/reference/src/replicann/arch/transformer.py:87-201: one encoder, one
decoder, one cross-decoder layer, no embedding, no stack, no head); SURVEY.md section 8(f).4 names the stacks as
the replication upstream would add next.  They are written here in the reference's own style
(arch/transformer.py:44-84: private `_`-prefixed sub-modules, keyword-only hyper-parameters, read-only
properties) out of the drop-in blocks of `arch/transformer.py`, so a stack's state_dict is the blocks' reference
state_dicts under `_blocks.<i>.`; everything runs on the library's kernels (tcgen05 GEMMs, fused attention,
LayerNorm, the embedding gather / deterministic scatter-add, fused softmax cross-entropy).

One deliberate addition: the reference's decoder blocks are NOT causal (their float `tril` buffer takes the
additive-mask branch), which is reproduced exactly by `causal=False`; a language model needs the
causal variant, so `causal=True` (the default of the decoder stack and of TransformerLM) swaps the buffer for the
boolean strictly-upper-triangular mask -- the reference's own `masked_fill(mask, -inf)` branch
(nn/attention.py:39-40).  The state_dict key (`_attn_mask`) and shape stay the same.
"""
import torch
import torch.nn as nn
from torch import Tensor

from replicann_b200.arch.transformer import TransformerCrossDecoder, TransformerDecoder, TransformerEncoder
from replicann_b200.nn.layers import Dropout, Embedding, Linear, cross_entropy


def _block_kwargs(ffn_bias, ffn_hidden_size, head_bias, proj_bias, p_dropout) -> dict:
    return dict(ffn_bias=ffn_bias, ffn_hidden_size=ffn_hidden_size, head_bias=head_bias, proj_bias=proj_bias,
                p_dropout=p_dropout)


def _make_causal(block: nn.Module, context_size: int) -> None:
    """True above the diagonal = masked out (nn/attention.py:39-40)."""
    block._attn_mask = torch.triu(torch.ones(context_size, context_size, dtype=torch.bool), diagonal=1)


class _Stack(nn.Module):
    @property
    def n_layers(self) -> int:
        return len(self._blocks)

    @property
    def embedding_size(self) -> int:
        return self._blocks[0].embedding_size

    @property
    def n_heads(self) -> int:
        return self._blocks[0].n_heads


class TransformerEncoderStack(_Stack):
    def __init__(
        self,
        n_layers: int,
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
        kw = _block_kwargs(ffn_bias, ffn_hidden_size, head_bias, proj_bias, p_dropout)
        self._blocks = nn.ModuleList(TransformerEncoder(n_heads, embedding_size, **kw) for _ in range(n_layers))

    def forward(self, x: Tensor, *, return_kv: bool = False) -> Tensor:
        """`return_kv=True`: the LAST block also returns its unprojected keys / values (arch/transformer.py:113-117),
        the memory a cross-decoder stack attends to."""
        for block in self._blocks[:-1]:
            x = block(x)
        return self._blocks[-1](x, return_kv=return_kv)


class TransformerDecoderStack(_Stack):
    def __init__(
        self,
        n_layers: int,
        n_heads: int,
        embedding_size: int,
        *,
        context_size: int,
        causal: bool = True,
        ffn_bias: bool = True,
        ffn_hidden_size: int | None = None,
        head_bias: bool = False,
        proj_bias: bool = True,
        p_dropout: float = 0.1,
    ) -> None:
        super().__init__()
        kw = _block_kwargs(ffn_bias, ffn_hidden_size, head_bias, proj_bias, p_dropout)
        self._blocks = nn.ModuleList(
            TransformerDecoder(n_heads, embedding_size, context_size=context_size, **kw) for _ in range(n_layers))
        self._causal = causal
        if causal:
            for block in self._blocks:
                _make_causal(block, context_size)

    @property
    def context_size(self) -> int:
        return self._blocks[0]._attn_mask.shape[0]

    @property
    def causal(self) -> bool:
        return self._causal

    def forward(self, x: Tensor) -> Tensor:
        for block in self._blocks:
            x = block(x)
        return x


class TransformerCrossDecoderStack(_Stack):
    def __init__(
        self,
        n_layers: int,
        n_heads: int,
        embedding_size: int,
        *,
        context_size: int,
        causal: bool = True,
        ffn_bias: bool = True,
        ffn_hidden_size: int | None = None,
        head_bias: bool = False,
        proj_bias: bool = True,
        p_dropout: float = 0.1,
    ) -> None:
        super().__init__()
        kw = _block_kwargs(ffn_bias, ffn_hidden_size, head_bias, proj_bias, p_dropout)
        self._blocks = nn.ModuleList(
            TransformerCrossDecoder(n_heads, embedding_size, context_size=context_size, **kw) for _ in range(n_layers))
        self._causal = causal
        if causal:
            for block in self._blocks:
                _make_causal(block, context_size)

    @property
    def context_size(self) -> int:
        return self._blocks[0]._attn_mask.shape[0]

    @property
    def causal(self) -> bool:
        return self._causal

    def forward(self, x: Tensor, /, encoder_k: Tensor, encoder_v: Tensor) -> Tensor:
        for block in self._blocks:
            x = block(x, encoder_k, encoder_v)
        return x


class TransformerEmbedding(nn.Module):
    """Token embedding + learned positional embedding + dropout: `dropout(tok[ids] + pos[:T])`, one gather kernel."""

    def __init__(self, vocab_size: int, embedding_size: int, *, context_size: int, p_dropout: float = 0.1) -> None:
        super().__init__()
        self._tokens = Embedding(vocab_size, embedding_size)
        self._positions = Embedding(context_size, embedding_size)
        self._dropout = Dropout(p_dropout)

    @property
    def vocab_size(self) -> int:
        return self._tokens.num_embeddings

    @property
    def context_size(self) -> int:
        return self._positions.num_embeddings

    @property
    def embedding_size(self) -> int:
        return self._tokens.embedding_dim

    @property
    def tokens(self) -> Embedding:
        return self._tokens

    def forward(self, ids: Tensor, /) -> Tensor:
        return self._dropout(self._tokens(ids, positions=self._positions))


class TransformerLM(nn.Module):
    """Decoder-only language model: embeddings -> N decoder blocks -> Linear head over the vocabulary.

    `forward(ids)` returns logits [B, T, V]; `forward(ids, targets)` returns the mean next-token cross-entropy
    (the fused softmax-CE kernel over [B*T, V], F.cross_entropy semantics).  `tie_weights=True` shares the head's
    matrix with the token table (the head then has no bias); the table's gradient is the sum of both uses."""

    def __init__(
        self,
        vocab_size: int,
        n_layers: int,
        n_heads: int,
        embedding_size: int,
        *,
        context_size: int,
        causal: bool = True,
        tie_weights: bool = False,
        head_bias: bool = False,
        ffn_hidden_size: int | None = None,
        p_dropout: float = 0.1,
    ) -> None:
        super().__init__()
        self._embedding = TransformerEmbedding(vocab_size, embedding_size, context_size=context_size,
                                               p_dropout=p_dropout)
        self._decoder = TransformerDecoderStack(n_layers, n_heads, embedding_size, context_size=context_size,
                                                causal=causal, ffn_hidden_size=ffn_hidden_size, head_bias=head_bias,
                                                p_dropout=p_dropout)
        self._head = Linear(embedding_size, vocab_size, bias=not tie_weights)
        self._tied = tie_weights
        if tie_weights:
            self._head.weight = self._embedding.tokens.weight

    @property
    def vocab_size(self) -> int:
        return self._embedding.vocab_size

    @property
    def context_size(self) -> int:
        return self._embedding.context_size

    @property
    def embedding_size(self) -> int:
        return self._embedding.embedding_size

    @property
    def n_layers(self) -> int:
        return self._decoder.n_layers

    @property
    def tied(self) -> bool:
        return self._tied

    def forward(self, ids: Tensor, /, targets: Tensor | None = None) -> Tensor:
        x = self._decoder(self._embedding(ids))
        logits = self._head(x)
        if targets is None:
            return logits
        return cross_entropy(logits.reshape(-1, logits.shape[-1]), targets.reshape(-1))
