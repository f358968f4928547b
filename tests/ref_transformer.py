"""Functional torch (CPU, autograd) restatement of the reference's attention / transformer-block semantics,
driven by a state_dict -- the live reference for GPU parity tests at sizes beyond the golden fixtures
(/root/reference is not available on the GPU box).  Mirrors oracle/transformer_numpy.py, which is pinned
against the reference's own outputs."""
import torch
import torch.nn.functional as F


def _lin(x, p, name):
    return F.linear(x, p[name + ".weight"], p.get(name + ".bias"))


def _attention(q, k, v, in_dim, mask):
    w = q @ k.transpose(-2, -1) * (1.0 / in_dim ** 0.5)
    if mask is not None:
        if mask.dtype == torch.bool:
            w = w.masked_fill(mask, float("-inf"))
        elif mask.dtype == torch.float32:
            w = w + mask
        else:
            raise TypeError(f"Unexpected type `{mask.dtype}` of the `mask`.")
    return F.softmax(w, dim=-1) @ v


def mhsa(p, pre, x, H, mask=None, return_kv=False):
    E = x.shape[-1]
    zs, ks, vs = [], [], []
    for h in range(H):
        q, k, v = (_lin(x, p, f"{pre}._heads.{h}._{n}") for n in ("query", "key", "value"))
        zs.append(_attention(q, k, v, E, mask)); ks.append(k); vs.append(v)
    z = torch.cat(zs, -1)
    if return_kv:
        return z, torch.cat(ks, -1), torch.cat(vs, -1)
    return _lin(z, p, f"{pre}._proj")


def mhca(p, pre, x, k, v, H, mask=None):
    E = x.shape[-1]
    sp = E // H
    zs = [_attention(_lin(x, p, f"{pre}._heads.{h}._query"), k[..., h * sp:(h + 1) * sp], v[..., h * sp:(h + 1) * sp], E, mask)
          for h in range(H)]
    return _lin(torch.cat(zs, -1), p, f"{pre}._proj")


def _ln(x, p, name):
    return F.layer_norm(x, (x.shape[-1],), p[name + ".weight"], p[name + ".bias"], 1e-5)


def _ffn(p, x):
    return _lin(F.relu(_lin(x, p, "_ffn._upscale")), p, "_ffn._downscale")


def encoder(p, x, H, return_kv=False):
    if not return_kv:
        x = _ln(x + mhsa(p, "_attn", x, H), p, "_attn_ln")
        return _ln(x + _ffn(p, x), p, "_ffn_ln")
    z, k, v = mhsa(p, "_attn", x, H, return_kv=True)
    x = _ln(x + z, p, "_attn_ln")
    return _ln(x + _ffn(p, x), p, "_ffn_ln"), k, v


def decoder(p, x, H):
    T = x.shape[1]
    x = _ln(x + mhsa(p, "_attn", x, H, p["_attn_mask"][:T, :T]), p, "_attn_ln")
    return _ln(x + _ffn(p, x), p, "_ffn_ln")


def cross_decoder(p, x, ek, ev, H):
    T = x.shape[1]
    x = _ln(x + mhsa(p, "_attn", x, H, p["_attn_mask"][:T, :T]), p, "_attn_ln")
    x = _ln(x + mhca(p, "_cross_attn", x, ek, ev, H), p, "_cross_attn_ln")
    return _ln(x + _ffn(p, x), p, "_ffn_ln")
