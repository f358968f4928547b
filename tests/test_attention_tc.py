"""Tensor-core attention pipeline (batched tcgen05 GEMMs + row softmax) against the fused fp32 kernel of the same
library run on fp32 copies of the inputs, and against plain torch on the CPU: forward and all three gradients,
bool and additive masks, cross attention (Tk != T), packed and separate projections, and dropout (both paths draw
the same Philox mask, so they must agree element for element up to bf16 rounding).
Reference semantics: nn/attention.py:28-50 (scale = 1/sqrt(E), masks, softmax, dropout, @ v)."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def rel(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return ((a - b).abs().max() / (b.abs().max() + 1e-12)).item()


def torch_ref(q, k, v, H, hs, scale, mask):
    """CPU fp32 reference on [B,T,H*hs] tensors; returns z and grads wrt q,k,v for dz = ones-ish."""
    B, T, _ = q.shape
    Tk = k.shape[1]
    qh = q.view(B, T, H, hs).transpose(1, 2)
    kh = k.view(B, Tk, H, hs).transpose(1, 2)
    vh = v.view(B, Tk, H, hs).transpose(1, 2)
    w = qh @ kh.transpose(-1, -2) * scale
    if mask is not None:
        w = w.masked_fill(mask, float("-inf")) if mask.dtype == torch.bool else w + mask
    w = torch.softmax(w, dim=-1)
    return (w @ vh).transpose(1, 2).reshape(B, T, H * hs)


@pytest.mark.parametrize("B,H,T,Tk,hs,mask_kind,packed", [
    (2, 4, 128, 128, 64, None, True),
    (2, 4, 128, 128, 64, "bool", False),
    (3, 2, 72, 200, 64, "float", False),      # cross attention, ragged against the 128 / 64 tiles
    (1, 2, 256, 256, 128, None, False),
    (8, 12, 512, 512, 64, "float", True),     # Track R shape L
])
def test_tc_attention_matches_torch(B, H, T, Tk, hs, mask_kind, packed):
    from replicann_b200 import functional as RF
    g = torch.Generator().manual_seed(B * 1000 + T)
    E = H * hs
    if packed:
        assert T == Tk
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
        mask = torch.tril(torch.ones(T, Tk))          # the reference decoder's additive float mask
    scale = 1.0 / math.sqrt(E)
    dz = (torch.randn(B, T, E, generator=g) * 0.1).to(torch.bfloat16)
    assert RF.attention_tc_ok(qd, kd, vd, H, hs)

    qr, kr, vr = (t.float().clone().requires_grad_(True) for t in (q, k, v))
    zr = torch_ref(qr, kr, vr, H, hs, scale, mask)
    zr.backward(dz.float())

    md = mask.cuda() if mask is not None else None
    z, saved = RF.attention_tc_fwd(qd, kd, vd, H, hs, scale, md)
    dq, dk, dv = (torch.empty_strided(t.shape, t.stride(), dtype=t.dtype, device="cuda") for t in (qd, kd, vd))
    RF.attention_tc_bwd(qd, kd, vd, saved, dz.cuda(), H, hs, scale, dq, dk, dv, md)
    torch.cuda.synchronize()
    assert rel(z, zr) < 1e-2
    assert rel(dq, qr.grad) < 2e-2
    assert rel(dk, kr.grad) < 2e-2
    assert rel(dv, vr.grad) < 2e-2


def test_tc_attention_dropout_matches_fused_kernel():
    """Same seed/offset -> same Philox keep mask in both paths; compare with the fused kernel on fp32 inputs."""
    from replicann_b200 import functional as RF
    B, H, T, hs = 2, 4, 128, 64
    E = H * hs
    g = torch.Generator().manual_seed(5)
    q, k, v = ((torch.randn(B, T, E, generator=g) * 0.5).to(torch.bfloat16).cuda() for _ in range(3))
    dz = (torch.randn(B, T, E, generator=g) * 0.1).to(torch.bfloat16).cuda()
    scale, p, seed, off = 1.0 / math.sqrt(E), 0.1, 1234, 7
    z, saved = RF.attention_tc_fwd(q, k, v, H, hs, scale, None, p, seed, off)
    dq, dk, dv = (torch.empty_like(t) for t in (q, k, v))
    RF.attention_tc_bwd(q, k, v, saved, dz, H, hs, scale, dq, dk, dv, None, p, seed, off)
    qf, kf, vf, dzf = (t.float() for t in (q, k, v, dz))
    zf, lse = RF.attention_fwd(qf, kf, vf, H, hs, scale, None, p, seed, off)
    dqf, dkf, dvf = (torch.empty_like(t) for t in (qf, kf, vf))
    RF.attention_bwd(qf, kf, vf, zf, dzf, lse, H, hs, scale, dqf, dkf, dvf, None, p, seed, off)
    torch.cuda.synchronize()
    assert rel(z, zf) < 1e-2
    assert rel(dq, dqf) < 2e-2 and rel(dk, dkf) < 2e-2 and rel(dv, dvf) < 2e-2
    # dropout really happened: about p of the probabilities are zero in the second half of the saved buffer
    n = B * H * T * T
    pd = saved.view(torch.bfloat16)[n:]
    frac = (pd == 0).float().mean().item()
    assert 0.07 < frac < 0.13


def test_tc_attention_rejects_unsupported_shapes():
    from replicann_b200 import functional as RF
    q = torch.zeros(1, 16, 2 * 32, dtype=torch.bfloat16, device="cuda")
    assert not RF.attention_tc_ok(q, q, q, 2, 32)           # head size 32: stays on the fused kernel
    q = torch.zeros(1, 12, 64, dtype=torch.bfloat16, device="cuda")
    assert not RF.attention_tc_ok(q, q, q, 1, 64)           # T = 12 is not a multiple of 8
    with pytest.raises(Exception):
        RF.attention_tc_fwd(q, q, q, 1, 64, 0.125)


@pytest.mark.parametrize("B,H,T,Tk,hs,mask_kind,p_drop", [
    (2, 4, 128, 128, 64, None, 0.0),
    (6, 2, 72, 200, 32, "float", 0.0),       # cross attention, head size 32, ragged against the tiles
    (2, 4, 128, 128, 64, "bool", 0.1),
    (8, 12, 512, 512, 64, "float", 0.0),     # Track R shape L in the reference's default dtype
])
def test_f32_tc_attention_is_fp32_accurate(B, H, T, Tk, hs, mask_kind, p_drop):
    """fp32 attention on the tensor cores (three-piece bf16 splits): forward and gradients against the fused fp32
    kernel of the library (same Philox mask when dropout is on) at the fp32 tolerance of the north star (1e-4),
    and against float64 torch on the CPU when there is no dropout."""
    from replicann_b200 import functional as RF
    g = torch.Generator().manual_seed(B * 100 + Tk)
    E = H * hs
    q = (torch.randn(B, T, E, generator=g) * 0.5)
    k = (torch.randn(B, Tk, E, generator=g) * 0.5)
    v = (torch.randn(B, Tk, E, generator=g) * 0.5)
    dz = (torch.randn(B, T, E, generator=g) * 0.1)
    mask = None
    if mask_kind == "bool":
        mask = torch.triu(torch.ones(T, Tk, dtype=torch.bool), diagonal=1)
    elif mask_kind == "float":
        mask = torch.tril(torch.ones(T, Tk))
    scale = 1.0 / math.sqrt(E)
    qd, kd, vd, dzd = q.cuda(), k.cuda(), v.cuda(), dz.cuda()
    md = mask.cuda() if mask is not None else None
    assert RF.attention_f32_tc_ok(qd, kd, vd, H, hs)
    z, saved = RF.attention_f32_tc_fwd(qd, kd, vd, H, hs, scale, md, p_drop, 99, 3)
    dq, dk, dv = (torch.empty_like(t) for t in (qd, kd, vd))
    RF.attention_f32_tc_bwd(qd, kd, vd, saved, dzd, H, hs, scale, dq, dk, dv, md, p_drop, 99, 3)
    zf, lse = RF.attention_fwd(qd, kd, vd, H, hs, scale, md, p_drop, 99, 3)
    dqf, dkf, dvf = (torch.empty_like(t) for t in (qd, kd, vd))
    RF.attention_bwd(qd, kd, vd, zf, dzd, lse, H, hs, scale, dqf, dkf, dvf, md, p_drop, 99, 3)
    torch.cuda.synchronize()
    for a, b, name in ((z, zf, "z"), (dq, dqf, "dq"), (dk, dkf, "dk"), (dv, dvf, "dv")):
        assert rel(a, b) < 1e-4, (name, rel(a, b))
    if p_drop == 0.0:
        qr, kr, vr = (t.double().clone().requires_grad_(True) for t in (q, k, v))
        zr = torch_ref(qr, kr, vr, H, hs, scale, mask.double() if mask_kind == "float" else mask)
        zr.backward(dz.double())
        assert rel(z, zr) < 1e-4 and rel(dq, qr.grad) < 1e-4 and rel(dk, kr.grad) < 1e-4 and rel(dv, vr.grad) < 1e-4
