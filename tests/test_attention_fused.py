"""Fused tcgen05 attention (rnn_attention_fused_*): scores and probabilities never leave the SM.

Checked against plain torch on the CPU (forward and all three gradients) for: packed and separate projections,
bool and additive float masks (incl. broadcast batch / row strides), cross attention with T != Tk, lengths that are
not multiples of the 128-row tiles, both supported head sizes, and the survey's Track-R shapes S and L; and against
the library's fp32 FMA kernel for dropout (same Philox counters -> same keep mask -> results agree up to bf16
rounding).  Reference semantics: nn/attention.py:28-50.  Tolerance: bf16 1e-2 forward, 2e-2 gradients (max-norm),
gradient cosine >= 0.999 (north_star).
The memory claim -- no [B,H,T,Tk] tensor -- is checked by the allocator: peak memory of a forward + backward pass
at shape L stays far below the size of one such tensor."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def rel(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return ((a - b).abs().max() / (b.abs().max() + 1e-12)).item()


def cosine(a, b):
    a, b = a.detach().double().cpu().flatten(), b.detach().double().cpu().flatten()
    return (a @ b / (a.norm() * b.norm() + 1e-300)).item()


def torch_ref(q, k, v, H, hs, scale, mask, keep=None):
    B, T, _ = q.shape
    Tk = k.shape[1]
    qh = q.view(B, T, H, hs).transpose(1, 2)
    kh = k.view(B, Tk, H, hs).transpose(1, 2)
    vh = v.view(B, Tk, H, hs).transpose(1, 2)
    w = qh @ kh.transpose(-1, -2) * scale
    if mask is not None:
        m = mask if mask.dim() < 3 else mask.unsqueeze(1)
        w = w.masked_fill(m, float("-inf")) if mask.dtype == torch.bool else w + m
    w = torch.softmax(w, dim=-1)
    return (w @ vh).transpose(1, 2).reshape(B, T, H * hs)


CASES = [
    # B, H, T, Tk, hs, mask, packed
    (2, 4, 128, 128, 64, None, True),
    (2, 4, 128, 128, 64, "bool", False),
    (3, 2, 72, 200, 64, "float", False),        # cross attention, ragged against the 128-row tiles
    (1, 2, 256, 256, 128, None, False),         # head size 128
    (2, 2, 264, 136, 128, "bool_b", False),     # head size 128, ragged, per-batch key-padding mask [B,1,Tk]
    (2, 3, 8, 8, 64, "float", True),            # a single partial tile
    (64, 8, 128, 128, 64, None, True),          # Track R shape S
    (8, 12, 512, 512, 64, "float", True),       # Track R shape L (the reference decoder's additive tril mask)
    (2, 2, 384, 640, 64, "bool_row", False),    # several key tiles, mask broadcast over rows [1,Tk]
]


@pytest.mark.parametrize("B,H,T,Tk,hs,mask_kind,packed", CASES)
def test_fused_attention_matches_torch(B, H, T, Tk, hs, mask_kind, packed):
    from replicann_b200 import functional as RF
    g = torch.Generator().manual_seed(B * 1000 + T + Tk)
    E = H * hs
    if packed:
        qkv = (torch.randn(B, T, 3 * E, generator=g) * 0.5).to(torch.bfloat16)
        q, k, v = qkv[..., :E], qkv[..., E:2 * E], qkv[..., 2 * E:]
        dev = qkv.cuda()
        qd, kd, vd = dev[..., :E], dev[..., E:2 * E], dev[..., 2 * E:]
    else:
        q = (torch.randn(B, T, E, generator=g) * 0.5).to(torch.bfloat16)
        k = (torch.randn(B, Tk, E, generator=g) * 0.5).to(torch.bfloat16)
        v = (torch.randn(B, Tk, E, generator=g) * 0.5).to(torch.bfloat16)
        qd, kd, vd = q.cuda(), k.cuda(), v.cuda()
    mask = None
    if mask_kind == "bool":
        mask = torch.triu(torch.ones(T, Tk, dtype=torch.bool), diagonal=1)
    elif mask_kind == "float":
        mask = torch.tril(torch.ones(T, Tk))
    elif mask_kind == "bool_b":
        mask = torch.rand(B, 1, Tk, generator=g) < 0.3
        mask[..., 0] = False
    elif mask_kind == "bool_row":
        mask = torch.rand(1, Tk, generator=g) < 0.5
        mask[..., 0] = False
    scale = 1.0 / math.sqrt(E)
    dz = (torch.randn(B, T, E, generator=g) * 0.1).to(torch.bfloat16)
    assert RF.attention_fused_ok(qd, kd, vd, H, hs)

    qr, kr, vr = (t.float().clone().requires_grad_(True) for t in (q, k, v))
    zr = torch_ref(qr, kr, vr, H, hs, scale, mask)
    zr.backward(dz.float())

    md = RF.broadcast_mask(mask.cuda(), qd, kd, vd, H, hs) if mask is not None else None
    z, lse = RF.attention_fused_fwd(qd, kd, vd, H, hs, scale, md)
    dq, dk, dv = (torch.empty_strided(t.shape, t.stride(), dtype=t.dtype, device="cuda") for t in (qd, kd, vd))
    RF.attention_fused_bwd(qd, kd, vd, z, dz.cuda(), lse, H, hs, scale, dq, dk, dv, md)
    torch.cuda.synchronize()
    print(f"STAT fused B{B} H{H} T{T} Tk{Tk} hs{hs} {mask_kind}: z {rel(z, zr):.2e} dq {rel(dq, qr.grad):.2e} "
          f"dk {rel(dk, kr.grad):.2e} dv {rel(dv, vr.grad):.2e}", flush=True)
    assert torch.isfinite(z).all()
    assert rel(z, zr) < 1e-2
    for got, ref, name in ((dq, qr.grad, "dq"), (dk, kr.grad, "dk"), (dv, vr.grad, "dv")):
        assert rel(got, ref) < 2e-2, name
        assert cosine(got, ref) >= 0.999, name
    # the saved log-sum-exp is the one of the masked, scaled scores
    w = (qr.view(B, T, H, hs).transpose(1, 2) @ kr.view(B, Tk, H, hs).transpose(1, 2).transpose(-1, -2)).detach() * scale
    if mask is not None:
        m = mask if mask.dim() < 3 else mask.unsqueeze(1)
        w = w.masked_fill(m, float("-inf")) if mask.dtype == torch.bool else w + m
    assert rel(lse, torch.logsumexp(w, dim=-1)) < 1e-3


@pytest.mark.parametrize("hs,T,Tk", [(64, 128, 128), (128, 136, 264)])
def test_fused_attention_dropout_matches_fp32_kernel(hs, T, Tk):
    """Same seed / offset -> the same Philox keep mask as the library's fp32 FMA kernel (and as the unfused
    tensor-core pipeline): outputs and gradients agree up to bf16 rounding, and about p of the mass is dropped."""
    from replicann_b200 import functional as RF
    B, H = 2, 2
    E = H * hs
    g = torch.Generator().manual_seed(5)
    q = (torch.randn(B, T, E, generator=g) * 0.5).to(torch.bfloat16).cuda()
    k, v = ((torch.randn(B, Tk, E, generator=g) * 0.5).to(torch.bfloat16).cuda() for _ in range(2))
    dz = (torch.randn(B, T, E, generator=g) * 0.1).to(torch.bfloat16).cuda()
    scale, p, seed, off = 1.0 / math.sqrt(E), 0.1, 1234, 7
    z, lse = RF.attention_fused_fwd(q, k, v, H, hs, scale, None, p, seed, off)
    dq, dk, dv = (torch.empty_like(t) for t in (q, k, v))
    RF.attention_fused_bwd(q, k, v, z, dz, lse, H, hs, scale, dq, dk, dv, None, p, seed, off)
    qf, kf, vf, dzf = (t.float() for t in (q, k, v, dz))
    zf, lsef = RF.attention_fwd(qf, kf, vf, H, hs, scale, None, p, seed, off)
    dqf, dkf, dvf = (torch.empty_like(t) for t in (qf, kf, vf))
    RF.attention_bwd(qf, kf, vf, zf, dzf, lsef, H, hs, scale, dqf, dkf, dvf, None, p, seed, off)
    z0, _ = RF.attention_fused_fwd(q, k, v, H, hs, scale, None, 0.0, seed, off)
    torch.cuda.synchronize()
    assert rel(z, zf) < 1e-2
    assert rel(dq, dqf) < 2e-2 and rel(dk, dkf) < 2e-2 and rel(dv, dvf) < 2e-2
    assert rel(z, z0) > 2e-2                      # dropout really changed the output
    assert rel(lse, lsef) < 1e-3                  # the log-sum-exp is that of the undropped probabilities


def test_attention_module_takes_the_fused_path_and_keeps_no_score_tensor():
    """MultiheadSelfAttention at Track-R shape L: forward + backward through the module, and the allocator's peak
    stays below what ONE [B,H,T,T] bf16 tensor would need on top of the inputs / outputs (the unfused pipeline
    keeps two of them plus fp32 scratch)."""
    from replicann_b200 import functional as RF
    from replicann_b200.nn._functions import AttentionFn
    B, H, T, hs = 8, 12, 512, 64
    E = H * hs
    g = torch.Generator().manual_seed(9)
    qkv = (torch.randn(B, T, 3 * E, generator=g) * 0.5).to(torch.bfloat16).cuda().requires_grad_(True)
    scale = 1.0 / math.sqrt(E)
    RF._workspaces.clear()
    torch.cuda.empty_cache()
    torch.cuda.synchronize()
    base = torch.cuda.memory_allocated()
    torch.cuda.reset_peak_memory_stats()
    z = AttentionFn.apply(qkv, None, None, None, scale, 0.0, H, hs, 0, 0, True)
    z.backward(torch.ones_like(z))
    torch.cuda.synchronize()
    peak = torch.cuda.max_memory_allocated() - base
    one_score_tensor = B * H * T * T * 2
    io = 3 * B * T * E * 2 * 2 + 2 * B * T * E * 2            # qkv + its gradient, z + its gradient
    assert peak < io + one_score_tensor // 2, (peak, io, one_score_tensor)
