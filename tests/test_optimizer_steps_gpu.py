"""Several optimizer steps in a row on the drop-in transformer blocks, against torch.optim on the CPU restatement
(tests/ref_transformer.py).  One step cannot show a gradient that was not cleared between steps; three can.

Covers: the packed per-head projections with gradients written in place into the flat bucket (optim.ParamPack),
the concatenating path used when no flat optimizer owns the parameters, `return_kv=True` (the output projection
receives no gradient), torch's default `model.zero_grad()` (set_to_none=True) between steps, and mask shapes that
broadcast the way the reference's `masked_fill` / `+=` do (nn/attention.py:39-42).
Tolerance: fp32 1e-4 (north_star)."""
import pytest
import torch

import ref_transformer as RT

pytestmark = pytest.mark.gpu


def rel(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return ((a - b).abs().max() / (b.abs().max() + 1e-12)).item()


def _twin(mod):
    """CPU leaf copies of a module's parameters / buffers for the functional reference."""
    p = {k: v.detach().float().cpu().clone() for k, v in mod.state_dict().items()}
    for k, _ in mod.named_parameters():
        p[k].requires_grad_(True)
    return p


@pytest.mark.parametrize("which,return_kv", [("adamw", False), ("sgd", False), ("adamw", True)])
def test_encoder_three_optimizer_steps_match_torch(which, return_kv):
    from replicann_b200.arch.transformer import TransformerEncoder
    from replicann_b200.optim import SGD, AdamW
    torch.manual_seed(0)
    H, E, B, T = 4, 64, 4, 24
    # .eval(): the heads keep their own attention dropout of 0.1 whatever the block is given (nn/attention.py:80-83)
    enc = TransformerEncoder(H, E, p_dropout=0.0, head_bias=True).cuda().eval()
    p = _twin(enc)
    names = [k for k, _ in enc.named_parameters()]
    wd = 0.0 if return_kv else 1e-2   # torch skips a parameter without gradient entirely; ours sees a zero gradient
    if which == "adamw":
        opt, t_opt = AdamW(enc.parameters(), lr=1e-3, weight_decay=wd), torch.optim.AdamW([p[k] for k in names], lr=1e-3, weight_decay=wd)
    else:
        opt = SGD(enc.parameters(), lr=0.05, momentum=0.9, weight_decay=wd)
        t_opt = torch.optim.SGD([p[k] for k in names], lr=0.05, momentum=0.9, weight_decay=wd)
    assert enc._attn._wpack.views() is not None and enc._attn._bpack.views() is not None
    g = torch.Generator().manual_seed(1)
    for step in range(3):
        x = torch.randn(B, T, E, generator=g)
        tgt = torch.randn(B, T, E, generator=g)
        # reference
        t_opt.zero_grad()
        yr = RT.encoder(p, x, H, return_kv=return_kv)
        yr = yr[0] + 0.1 * yr[1] + 0.2 * yr[2] if return_kv else yr
        ((yr - tgt) ** 2).mean().backward()
        gref = {k: (p[k].grad.clone() if p[k].grad is not None else None) for k in names}
        t_opt.step()
        # ours; alternate between the optimizer's zero_grad and torch's module-level one (set_to_none=True)
        if step == 1:
            enc.zero_grad()
        opt.zero_grad()
        yo = enc(x.cuda(), return_kv=return_kv)
        yo = yo[0] + 0.1 * yo[1] + 0.2 * yo[2] if return_kv else yo
        ((yo - tgt.cuda()) ** 2).mean().backward()
        # some gradients are analytically zero (a key bias shifts every score of a row alike and softmax does not see
        # it): compare those against the scale of the whole gradient, not against their own rounding noise
        gmax = max(float(v.abs().max()) for v in gref.values() if v is not None)
        for k, prm in enc.named_parameters():
            if gref[k] is None:
                assert float(prm.grad.abs().max()) == 0.0, (step, k)   # cleared, not stale
            else:
                err = float((prm.grad.detach().float().cpu() - gref[k]).abs().max())
                assert err < 2e-4 * max(float(gref[k].abs().max()), 1e-3 * gmax), (step, k, err)
        opt.step()
        for k, prm in enc.named_parameters():
            assert rel(prm, p[k]) < 1e-4, (step, k, rel(prm, p[k]))


def test_packed_projection_without_flat_optimizer_accumulates_like_autograd():
    """No flat optimizer: the projection matrix is concatenated per call and gradients are scattered into each
    head's own .grad; two backward passes without clearing must ADD (plain autograd semantics)."""
    from replicann_b200.nn.attention import MultiheadSelfAttention
    torch.manual_seed(2)
    H, hs, E = 2, 16, 32
    att = MultiheadSelfAttention(H, hs, E, p_dropout=0.0).cuda().eval()
    assert att._wpack.views() is None
    p = _twin(att)
    x = torch.randn(3, 8, E, generator=torch.Generator().manual_seed(3))
    # reference through the same functional code, with the key prefix of a bare module
    pp = {"_m." + k: v for k, v in p.items()}
    for _ in range(2):
        RT.mhsa(pp, "_m", x, H).square().sum().backward()
        att(x.cuda()).square().sum().backward()
    for k, prm in att.named_parameters():
        assert rel(prm.grad, pp["_m." + k].grad) < 2e-4, k


@pytest.mark.parametrize("shape_kind", ["Tk", "1xTk", "Bx1xTk", "Tx1", "BxTxTk"])
@pytest.mark.parametrize("dtype,mask_dtype", [(torch.float32, torch.bool), (torch.float32, torch.float32),
                                              (torch.bfloat16, torch.bool)])
def test_masks_broadcast_like_the_reference(shape_kind, dtype, mask_dtype):
    from replicann_b200.nn._functions import AttentionFn
    B, H, T, Tk, hs = 3, 2, 16, 24, 64
    g = torch.Generator().manual_seed(4)
    q = torch.randn(B, T, H * hs, generator=g).to(dtype)
    k = torch.randn(B, Tk, H * hs, generator=g).to(dtype)
    v = torch.randn(B, Tk, H * hs, generator=g).to(dtype)
    shape = {"Tk": (Tk,), "1xTk": (1, Tk), "Bx1xTk": (B, 1, Tk), "Tx1": (T, 1), "BxTxTk": (B, T, Tk)}[shape_kind]
    if mask_dtype == torch.bool:
        mask = torch.rand(shape, generator=g) < 0.3
        mask[..., 0] = False     # never a fully masked row (softmax of all -inf is NaN in the reference too)
    else:
        mask = torch.randn(shape, generator=g)
    scale = 1.0 / (H * hs) ** 0.5
    qf, kf, vf = (t.float().reshape(B, -1, H, hs).transpose(1, 2) for t in (q, k, v))
    w = qf @ kf.transpose(-2, -1) * scale                       # [B, H, T, Tk]
    mb = mask if mask.dim() < 3 else mask.unsqueeze(1)          # the reference applies it per head on [B, T, Tk]
    w = w.masked_fill(mb, float("-inf")) if mask_dtype == torch.bool else w + mb
    ref = (torch.softmax(w, -1) @ vf).transpose(1, 2).reshape(B, T, H * hs)
    got = AttentionFn.apply(q.cuda(), k.cuda(), v.cuda(), mask.cuda(), scale, 0.0, H, hs, 0, 0, False)
    assert rel(got, ref) < (1e-4 if dtype == torch.float32 else 1e-2)


def test_bad_mask_and_operand_shapes_are_refused():
    from replicann_b200.nn._functions import AttentionFn
    B, H, T, Tk, hs = 2, 2, 8, 16, 8
    q = torch.randn(B, T, H * hs).cuda()
    k = torch.randn(B, Tk, H * hs).cuda()
    v = torch.randn(B, Tk, H * hs).cuda()
    run = lambda q_, k_, v_, m: AttentionFn.apply(q_, k_, v_, m, 0.25, 0.0, H, hs, 0, 0, False)   # noqa: E731
    for bad in [(Tk + 1,), (T, Tk + 1), (B + 1, T, Tk), (5, Tk), (1, 1, T, Tk)]:
        with pytest.raises(RuntimeError):
            run(q, k, v, torch.zeros(bad, dtype=torch.bool).cuda())
    with pytest.raises(ValueError):
        run(q, k[..., :-8], v, None)                 # k narrower than n_heads * head_size
    with pytest.raises(ValueError):
        run(q, k[:1], v[:1], None)                   # batch mismatch
    with pytest.raises(ValueError):
        run(q, k, v[:, :-1], None)                   # k / v lengths differ
