"""GPU parity of the encoder / decoder stacks, embeddings and the language-model loss.

(a) fp32 against golden logits / loss / gradients of a torch-CPU network assembled from the reference's OWN block
    classes (tests/golden/transformer_stacks.npz, make_goldens.py `transformer_stacks`): 1e-4 (north_star);
(b) bf16 at a realistic shape against a live torch-CPU restatement on the same bf16-rounded weights: logits / loss
    within 1e-2 and gradient cosine >= 0.999;
(c) the embedding kernels alone: bit-exact gather, deterministic scatter-add against index_add_ in float64.
"""
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.nn.functional as F

import ref_transformer as RT

pytestmark = pytest.mark.gpu
G = Path(__file__).resolve().parent / "golden"


def rel(a, b):
    a, b = a.detach().float().cpu(), torch.as_tensor(b).detach().float().cpu()
    return ((a - b).abs().max() / (b.abs().max() + 1e-12)).item()


def cosine(a, b):
    a, b = a.detach().double().cpu().flatten(), torch.as_tensor(b).detach().double().cpu().flatten()
    return (a @ b / (a.norm() * b.norm() + 1e-300)).item()


@pytest.fixture(scope="module")
def ts():
    return np.load(G / "transformer_stacks.npz")


def _sd(ts, tag):
    pre = f"{tag}:sd:"
    return {k[len(pre):]: torch.from_numpy(ts[k]) for k in ts.files if k.startswith(pre)}


def _check_grads(mod, ts, tag, tol=2e-4):
    n_checked = 0
    for n, p in mod.named_parameters():
        key = f"{tag}:grad:{n}"
        if key in ts.files:
            assert p.grad is not None, n
            assert rel(p.grad, ts[key]) < tol, (n, rel(p.grad, ts[key]))
            n_checked += 1
    assert n_checked >= 10


@pytest.mark.parametrize("tag", ["lm", "lm_tied_noncausal"])
@pytest.mark.parametrize("with_optimizer", [False, True])
def test_language_model_golden_fp32(ts, tag, with_optimizer):
    from replicann_b200.arch.stacks import TransformerLM
    from replicann_b200.optim import SGD
    V, L, H, E, ctx, causal, tie = (int(v) for v in ts[f"{tag}:cfg"])
    lm = TransformerLM(V, L, H, E, context_size=ctx, causal=bool(causal), tie_weights=bool(tie), p_dropout=0.0)
    res = lm.load_state_dict(_sd(ts, tag), strict=True)
    assert not res.missing_keys and not res.unexpected_keys
    lm = lm.cuda().eval()
    if with_optimizer:   # gradients then land in the flat buckets (packed projections, tied table accumulation)
        opt = SGD(lm.parameters(), lr=0.0)
        opt.zero_grad()
    ids = torch.from_numpy(ts[f"{tag}:ids"]).cuda()
    tgt = torch.from_numpy(ts[f"{tag}:targets"]).cuda()
    logits = lm(ids)
    assert tuple(logits.shape) == (ids.shape[0], ids.shape[1], V)
    assert rel(logits, ts[f"{tag}:logits"]) < 1e-4
    loss = lm(ids, tgt)
    assert abs(loss.item() - float(ts[f"{tag}:loss"])) < 1e-4 * max(1.0, abs(float(ts[f"{tag}:loss"])))
    loss.backward()
    _check_grads(lm, ts, tag)
    if tie:
        assert lm._head.weight is lm._embedding.tokens.weight


def test_encoder_and_cross_decoder_stack_golden_fp32(ts):
    from replicann_b200.arch.stacks import TransformerCrossDecoderStack, TransformerEncoderStack

    def wsum(o):
        return (o * torch.arange(o.numel(), dtype=torch.float32, device=o.device).reshape(o.shape).cos()).sum()

    enc = TransformerEncoderStack(2, 2, 16, p_dropout=0.0)
    enc.load_state_dict(_sd(ts, "encstack"), strict=True)
    enc = enc.cuda().eval()
    x = torch.from_numpy(ts["encstack:x"]).cuda().requires_grad_()
    y, k, v = enc(x, return_kv=True)
    for got, key in ((y, "y"), (k, "k"), (v, "v")):
        assert rel(got, ts[f"encstack:{key}"]) < 1e-4, key
    (wsum(y) + 0.5 * wsum(k) - 0.25 * wsum(v)).backward()
    assert rel(x.grad, ts["encstack:dx"]) < 2e-4
    _check_grads(enc, ts, "encstack")

    xd = TransformerCrossDecoderStack(2, 2, 16, context_size=12, p_dropout=0.0)
    xd.load_state_dict(_sd(ts, "xdecstack"), strict=True)
    xd = xd.cuda().eval()
    assert xd._blocks[0]._attn_mask.dtype == torch.bool
    x2 = torch.from_numpy(ts["xdecstack:x"]).cuda().requires_grad_()
    ek = torch.from_numpy(ts["encstack:k"]).cuda().requires_grad_()
    ev = torch.from_numpy(ts["encstack:v"]).cuda().requires_grad_()
    out = xd(x2, ek, ev)
    assert rel(out, ts["xdecstack:y"]) < 1e-4
    wsum(out).backward()
    assert rel(x2.grad, ts["xdecstack:dx"]) < 2e-4
    assert rel(ek.grad, ts["xdecstack:dek"]) < 2e-4 and rel(ev.grad, ts["xdecstack:dev"]) < 2e-4
    _check_grads(xd, ts, "xdecstack")


def _ref_lm(p, ids, H, L, tgt):
    """torch-CPU restatement of TransformerLM on a state_dict (tests/ref_transformer.py blocks)."""
    T = ids.shape[1]
    x = p["_embedding._tokens.weight"][ids] + p["_embedding._positions.weight"][:T]
    for i in range(L):
        pre = f"_decoder._blocks.{i}."
        x = RT.decoder({k[len(pre):]: v for k, v in p.items() if k.startswith(pre)}, x, H)
    logits = F.linear(x, p["_head.weight"], p.get("_head.bias"))
    return logits, F.cross_entropy(logits.reshape(-1, logits.shape[-1]), tgt.reshape(-1))


@pytest.mark.parametrize("tie", [False, True])
def test_language_model_bf16_training_step(tie):
    """bf16, V=4096, 2 layers of E=256 / H=4 (hs=64: the tensor-core attention path), T=128, causal."""
    from replicann_b200.arch.stacks import TransformerLM
    from replicann_b200.optim import AdamW
    torch.manual_seed(0)
    V, L, H, E, ctx, B, T = 4096, 2, 4, 256, 128, 4, 128
    lm = TransformerLM(V, L, H, E, context_size=ctx, tie_weights=tie, p_dropout=0.0).cuda().to(torch.bfloat16).eval()
    with torch.no_grad():   # embeddings at N(0,1) make the tied logits huge; a trained-model-like scale keeps CE sane
        lm._embedding.tokens.weight.mul_(0.05)
        lm._embedding._positions.weight.mul_(0.05)
    p = {k: (v.detach().float() if v.is_floating_point() else v.detach()).cpu().clone()
         for k, v in lm.state_dict().items()}
    for k, _ in lm.named_parameters():
        p[k].requires_grad_(True)
    if tie:
        p["_head.weight"] = p["_embedding._tokens.weight"]
    g = torch.Generator().manual_seed(1)
    ids = torch.randint(0, V, (B, T), generator=g)
    ids[:, ::7] = 11                                   # a frequent token: many rows scatter into one table row
    tgt = torch.randint(0, V, (B, T), generator=g)
    logits_r, loss_r = _ref_lm(p, ids, H, L, tgt)
    loss_r.backward()
    opt = AdamW(lm.parameters(), lr=1e-3)
    opt.zero_grad()
    logits = lm(ids.cuda())
    loss = lm(ids.cuda(), tgt.cuda())
    loss.backward()
    assert rel(logits, logits_r) < 1e-2
    assert abs(loss.item() - loss_r.item()) < 1e-2 * max(1.0, abs(loss_r.item()))
    names = [k for k, _ in lm.named_parameters()]
    allo = torch.cat([prm.grad.float().flatten().cpu() for _, prm in lm.named_parameters()])
    allr = torch.cat([p[k].grad.flatten() for k in names])
    assert cosine(allo, allr) >= 0.999, cosine(allo, allr)
    worst = min((cosine(prm.grad, p[k].grad), k) for k, prm in lm.named_parameters())
    print("STAT lm bf16 tie", tie, "logits", rel(logits, logits_r), "global cos", cosine(allo, allr), "worst", worst,
          flush=True)
    assert worst[0] >= 0.99, worst
    opt.step()
    assert torch.isfinite(lm(ids.cuda(), tgt.cuda())).item()


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("B,T,V,E", [(3, 17, 50, 64), (8, 512, 1000, 768), (2, 33, 7, 8)])
def test_embedding_kernels(dtype, B, T, V, E):
    from replicann_b200 import functional as RF
    g = torch.Generator().manual_seed(B * T + V)
    tok = torch.randn(V, E, generator=g).to(dtype)
    pos = torch.randn(T + 3, E, generator=g).to(dtype)
    ids = torch.randint(0, V, (B, T), generator=g)
    ids[0, : min(T, 9)] = V - 1
    out = RF.embedding_fwd(ids.cuda(), tok.cuda(), pos.cuda()[:T].contiguous())
    ref = (tok.float()[ids] + pos.float()[:T]).to(dtype)
    assert torch.equal(out.cpu(), ref)                                   # one add, one rounding: bit-exact
    assert torch.equal(RF.embedding_fwd(ids.cuda(), tok.cuda()).cpu(), tok[ids])
    dy = torch.randn(B, T, E, generator=g).to(dtype)
    dtok = torch.full((V, E), 7.0, dtype=dtype).cuda()                   # overwrite mode must not see the 7s
    dpos = torch.full((T, E), 7.0, dtype=dtype).cuda()
    RF.embedding_bwd(ids.cuda(), dy.cuda(), dtok=dtok, dpos=dpos)
    ref_tok = torch.zeros(V, E, dtype=torch.float64).index_add_(0, ids.flatten(), dy.double().reshape(-1, E))
    ref_pos = dy.double().sum(0)
    tol = 1e-5 if dtype == torch.float32 else 1e-2
    assert rel(dtok, ref_tok) < tol and rel(dpos, ref_pos) < tol
    # run-to-run bitwise reproducible, and accumulate mode adds to what is there
    again = torch.empty_like(dtok)
    RF.embedding_bwd(ids.cuda(), dy.cuda(), dtok=again)
    assert torch.equal(again, dtok)
    RF.embedding_bwd(ids.cuda(), dy.cuda(), dtok=dtok, dpos=dpos, accumulate_tok=True, accumulate_pos=True)
    assert rel(dtok, 2 * ref_tok) < 2 * tol and rel(dpos, 2 * ref_pos) < 2 * tol
