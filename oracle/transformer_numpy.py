"""ORACLE -- TEST INFRASTRUCTURE ONLY (rules in oracle/cnn_numpy.py).

numpy restatement of the reference's attention layers and transformer blocks, forward pass, following
/root/reference/src/replicann line by line.  Parameters come in as a dict keyed like the reference's
state_dict.  PINNING: checked in tests/test_oracle.py against tests/golden/transformer_blocks.npz, which
tests/golden/make_goldens.py produced by running the reference's own classes (fp32, dropout off).
"""
from __future__ import annotations

import numpy as np


def linear(x, w, b=None):
    y = x @ w.T
    return y + b if b is not None else y


def layer_norm(x, g, b, eps=1e-5):
    """nn.LayerNorm(E) with torch defaults (arch/transformer.py:65,72,188)."""
    mu = x.mean(axis=-1, keepdims=True)
    var = x.var(axis=-1, keepdims=True)
    return (x - mu) / np.sqrt(var + eps) * g + b


def softmax(w):
    m = w.max(axis=-1, keepdims=True)
    m = np.where(np.isfinite(m), m, 0.0)
    e = np.exp(w - m)
    return e / e.sum(axis=-1, keepdims=True)


def attention(q, k, v, in_dim, mask=None):
    """_AttentionHead._attention (nn/attention.py:28-50): scale is 1/sqrt(in_dim) -- the EMBEDDING size, not
    the head size (:24); bool mask: True = masked out (-inf) (:39-40); float32 mask: additive (:41-42);
    any other dtype: TypeError (:43-44).  Dropout (:46) is the identity here (p = 0 / eval)."""
    w = q @ np.swapaxes(k, -2, -1) * (1.0 / np.sqrt(in_dim))
    if mask is not None:
        if mask.dtype == np.bool_:
            w = np.where(mask, -np.inf, w)
        elif mask.dtype == np.float32:
            w = w + mask
        else:
            raise TypeError(f"Unexpected type `{mask.dtype}` of the `mask`.")
    return softmax(w) @ v


def multihead_self_attention(p, pre, x, n_heads, mask=None, return_kv=False):
    """MultiheadSelfAttention.forward (nn/attention.py:175-192).  return_kv=True returns the concatenated
    (z, k, v) WITHOUT the output projection (:183-192)."""
    E = x.shape[-1]
    zs, ks, vs = [], [], []
    for h in range(n_heads):
        q = linear(x, p[f"{pre}._heads.{h}._query.weight"], p.get(f"{pre}._heads.{h}._query.bias"))
        k = linear(x, p[f"{pre}._heads.{h}._key.weight"], p.get(f"{pre}._heads.{h}._key.bias"))
        v = linear(x, p[f"{pre}._heads.{h}._value.weight"], p.get(f"{pre}._heads.{h}._value.bias"))
        zs.append(attention(q, k, v, E, mask)); ks.append(k); vs.append(v)
    z = np.concatenate(zs, axis=-1)
    if return_kv:
        return z, np.concatenate(ks, axis=-1), np.concatenate(vs, axis=-1)
    return linear(z, p[f"{pre}._proj.weight"], p.get(f"{pre}._proj.bias"))


def multihead_cross_attention(p, pre, x, k, v, n_heads, mask=None):
    """MultiheadCrossAttention.forward (nn/attention.py:149-162): k and v are split into chunks of width
    E // H (:153-154) and used unprojected; only the query is projected (:112)."""
    E = x.shape[-1]
    sp = E // n_heads
    zs = []
    for h in range(n_heads):
        q = linear(x, p[f"{pre}._heads.{h}._query.weight"], p.get(f"{pre}._heads.{h}._query.bias"))
        zs.append(attention(q, k[..., h * sp:(h + 1) * sp], v[..., h * sp:(h + 1) * sp], E, mask))
    return linear(np.concatenate(zs, axis=-1), p[f"{pre}._proj.weight"], p.get(f"{pre}._proj.bias"))


def ffn(p, pre, x):
    """_TransformerFFN.forward (arch/transformer.py:28-33)."""
    h = np.maximum(linear(x, p[f"{pre}._upscale.weight"], p.get(f"{pre}._upscale.bias")), 0)
    return linear(h, p[f"{pre}._downscale.weight"], p.get(f"{pre}._downscale.bias"))


def encoder(p, x, n_heads, return_kv=False):
    """TransformerEncoder.forward (arch/transformer.py:109-117)."""
    if not return_kv:
        x = layer_norm(x + multihead_self_attention(p, "_attn", x, n_heads), p["_attn_ln.weight"], p["_attn_ln.bias"])
        return layer_norm(x + ffn(p, "_ffn", x), p["_ffn_ln.weight"], p["_ffn_ln.bias"])
    z, k, v = multihead_self_attention(p, "_attn", x, n_heads, return_kv=True)
    x = layer_norm(x + z, p["_attn_ln.weight"], p["_attn_ln.bias"])
    return layer_norm(x + ffn(p, "_ffn", x), p["_ffn_ln.weight"], p["_ffn_ln.bias"]), k, v


def _block_mask(p, T):
    """The decoders' `_attn_mask` buffer as the reference feeds it to attention (arch/transformer.py:148): float32
    tril(ones) in the reference itself (additive branch); a bool buffer (True = masked out) in the causal stacks of
    replicann_b200.arch.stacks takes the masked_fill branch of nn/attention.py:39-40."""
    m = p["_attn_mask"][:T, :T]
    return m if m.dtype == np.bool_ else m.astype(np.float32)


def decoder(p, x, n_heads):
    """TransformerDecoder.forward (arch/transformer.py:147-151): the float32 tril(ones) buffer takes the
    ADDITIVE mask branch -- +1 on and below the diagonal -- so the block is not causal."""
    T = x.shape[1]
    mask = _block_mask(p, T)
    x = layer_norm(x + multihead_self_attention(p, "_attn", x, n_heads, mask), p["_attn_ln.weight"], p["_attn_ln.bias"])
    return layer_norm(x + ffn(p, "_ffn", x), p["_ffn_ln.weight"], p["_ffn_ln.bias"])


def cross_decoder(p, x, enc_k, enc_v, n_heads):
    """TransformerCrossDecoder.forward (arch/transformer.py:190-201)."""
    T = x.shape[1]
    mask = _block_mask(p, T)
    x = layer_norm(x + multihead_self_attention(p, "_attn", x, n_heads, mask), p["_attn_ln.weight"], p["_attn_ln.bias"])
    x = layer_norm(x + multihead_cross_attention(p, "_cross_attn", x, enc_k, enc_v, n_heads),
                   p["_cross_attn_ln.weight"], p["_cross_attn_ln.bias"])
    return layer_norm(x + ffn(p, "_ffn", x), p["_ffn_ln.weight"], p["_ffn_ln.bias"])


# ---------------------------------------------------------------------------------------------- stacks
# Not in the reference: compositions of its blocks (arch/transformer.py:87-201) with torch.nn.Embedding /
# nn.Linear / F.cross_entropy semantics.  PINNING: tests/golden/transformer_stacks.npz, produced by
# make_goldens.py from a torch-CPU network assembled out of the reference's own block classes.
def _sub(p, prefix):
    return {k[len(prefix):]: v for k, v in p.items() if k.startswith(prefix)}


def _n_blocks(p, prefix):
    return 1 + max(int(k[len(prefix):].split(".")[0]) for k in p if k.startswith(prefix))


def encoder_stack(p, x, n_heads, return_kv=False):
    n = _n_blocks(p, "_blocks.")
    for i in range(n - 1):
        x = encoder(_sub(p, f"_blocks.{i}."), x, n_heads)
    return encoder(_sub(p, f"_blocks.{n - 1}."), x, n_heads, return_kv=return_kv)


def decoder_stack(p, x, n_heads):
    for i in range(_n_blocks(p, "_blocks.")):
        x = decoder(_sub(p, f"_blocks.{i}."), x, n_heads)
    return x


def cross_decoder_stack(p, x, enc_k, enc_v, n_heads):
    for i in range(_n_blocks(p, "_blocks.")):
        x = cross_decoder(_sub(p, f"_blocks.{i}."), x, enc_k, enc_v, n_heads)
    return x


def embedding(p, ids):
    """tok[ids] + pos[arange(T)] (torch.nn.Embedding lookups)."""
    return p["_tokens.weight"][ids] + p["_positions.weight"][np.arange(ids.shape[1])]


def cross_entropy(logits, targets):
    """F.cross_entropy(logits [N, V], targets [N]) with mean reduction."""
    m = logits.max(axis=-1, keepdims=True)
    lse = m[:, 0] + np.log(np.exp(logits - m).sum(axis=-1))
    return float((lse - logits[np.arange(len(targets)), targets]).mean())


def language_model(p, ids, n_heads, targets=None):
    """embeddings -> decoder stack -> Linear head; with targets, the mean next-token cross-entropy."""
    x = embedding(_sub(p, "_embedding."), ids)
    x = decoder_stack(_sub(p, "_decoder."), x, n_heads)
    logits = linear(x, p["_head.weight"], p.get("_head.bias"))
    if targets is None:
        return logits
    return logits, cross_entropy(logits.reshape(-1, logits.shape[-1]), targets.reshape(-1))
