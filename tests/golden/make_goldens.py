"""Generate the golden fixtures under tests/golden/ (run in the build container, where /root/reference and
torch CPU are available):   python tests/golden/make_goldens.py

Track N (CNN operators / training steps): outputs of `torch` CPU -- the third-party dependency in which the
reference's arithmetic lives (pyproject.toml:11; locked 2.4.0, installed version recorded in each file).
Track R (transformer blocks): outputs of the reference's OWN classes imported from /root/reference/src.
The fixtures are small on purpose; they pin the oracle (tests/test_oracle.py) and the CUDA path
(tests/test_transformer_gpu.py and tests/test_oracle.py) without needing /root/reference at test time.
"""
import sys
from pathlib import Path

import numpy as np
import torch
import torch.nn.functional as F

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE.parent))
import ref_models as R  # noqa: E402

torch.set_num_threads(1)
META = {"torch_version": np.array(torch.__version__)}


def t(a):
    return torch.tensor(a)


def cnn_ops():
    rng = np.random.default_rng(20240601)
    out = dict(META)
    # ---- conv2d: (stride, pad) variants incl. 1x1/s2 and 7x7/s2
    for tag, (N, C, H, W, K, Rr, stride, pad) in {
        "c3s1": (2, 4, 9, 9, 6, 3, 1, 1), "c3s2": (2, 4, 10, 10, 6, 3, 2, 1), "c1s2": (2, 8, 8, 8, 4, 1, 2, 0),
        "c7s2": (1, 3, 20, 20, 4, 7, 2, 3), "c5p2": (2, 1, 12, 12, 6, 5, 1, 2)}.items():
        x = rng.standard_normal((N, C, H, W)).astype(np.float32)
        w = (rng.standard_normal((K, C, Rr, Rr)) * 0.3).astype(np.float32)
        b = rng.standard_normal(K).astype(np.float32)
        xt, wt, bt = t(x).requires_grad_(), t(w).requires_grad_(), t(b).requires_grad_()
        y = F.conv2d(xt, wt, bt, stride=stride, padding=pad)
        dy = rng.standard_normal(tuple(y.shape)).astype(np.float32)
        y.backward(t(dy))
        out.update({f"conv_{tag}_{k}": v for k, v in dict(
            x=x, w=w, b=b, sp=np.array([stride, pad]), y=y.detach().numpy(), dy=dy, dx=xt.grad.numpy(),
            dw=wt.grad.numpy(), db=bt.grad.numpy()).items()})
    # ---- batchnorm (training) with ReLU + residual
    x = (rng.standard_normal((4, 8, 5, 5)) * 2 + 0.5).astype(np.float32)
    g, b = (rng.random(8) + 0.5).astype(np.float32), rng.standard_normal(8).astype(np.float32)
    res = rng.standard_normal((4, 8, 5, 5)).astype(np.float32)
    dy = rng.standard_normal((4, 8, 5, 5)).astype(np.float32)
    xt, gt, bt, rt = t(x).requires_grad_(), t(g).requires_grad_(), t(b).requires_grad_(), t(res).requires_grad_()
    rm, rv = torch.zeros(8), torch.ones(8)
    y = F.relu(F.batch_norm(xt, rm, rv, gt, bt, True, 0.1, 1e-5) + rt)
    y.backward(t(dy))
    out.update({f"bn_{k}": v for k, v in dict(x=x, g=g, b=b, res=res, dy=dy, y=y.detach().numpy(), dx=xt.grad.numpy(),
                                              dg=gt.grad.numpy(), db=bt.grad.numpy(), dres=rt.grad.numpy(),
                                              rm=rm.numpy(), rv=rv.numpy()).items()})
    # ---- max pooling with ties and a NaN
    x = (np.round(rng.standard_normal((2, 3, 9, 9)) * 2) / 2).astype(np.float32)
    x[0, 1, 4, 4] = np.nan
    for tag, (k, s, p) in {"k3s2p1": (3, 2, 1), "k2s2": (2, 2, 0), "k3s1p1": (3, 1, 1)}.items():
        y, idx = F.max_pool2d(t(x), k, s, p, return_indices=True)
        out.update({f"maxpool_{tag}_x": x, f"maxpool_{tag}_ksp": np.array([k, s, p]), f"maxpool_{tag}_y": y.numpy(),
                    f"maxpool_{tag}_idx": idx.numpy()})
    # ---- avg pooling
    x = rng.standard_normal((2, 3, 8, 8)).astype(np.float32)
    out["avgpool_x"] = x
    out["avgpool_k2"] = F.avg_pool2d(t(x), 2).numpy()
    out["avgpool_k3s2p1"] = F.avg_pool2d(t(x), 3, 2, 1).numpy()
    out["avgpool_global"] = F.adaptive_avg_pool2d(t(x), 1).numpy()
    # ---- linear, cross entropy
    x, w, b = rng.standard_normal((5, 7)).astype(np.float32), rng.standard_normal((4, 7)).astype(np.float32), \
        rng.standard_normal(4).astype(np.float32)
    tg = rng.integers(0, 4, 5)
    xt, wt, bt = t(x).requires_grad_(), t(w).requires_grad_(), t(b).requires_grad_()
    lg = F.linear(xt, wt, bt)
    loss = F.cross_entropy(lg, t(tg))
    loss.backward()
    out.update({f"lin_{k}": v for k, v in dict(x=x, w=w, b=b, t=tg, logits=lg.detach().numpy(), loss=np.float32(loss.item()),
                                               dx=xt.grad.numpy(), dw=wt.grad.numpy(), db=bt.grad.numpy()).items()})
    # ---- optimizers: three steps each
    w0 = rng.standard_normal(64).astype(np.float32)
    gs = rng.standard_normal((3, 64)).astype(np.float32)
    for name, mk in {"sgd": lambda p: torch.optim.SGD([p], lr=0.1, momentum=0.9, weight_decay=1e-4),
                     "sgd_nesterov": lambda p: torch.optim.SGD([p], lr=0.1, momentum=0.9, nesterov=True),
                     "adamw": lambda p: torch.optim.AdamW([p], lr=1e-2, weight_decay=1e-2)}.items():
        p = torch.nn.Parameter(t(w0.copy()))
        opt = mk(p)
        for i in range(3):
            p.grad = t(gs[i].copy())
            opt.step()
        out[f"opt_{name}_w"] = p.detach().numpy()
    out["opt_w0"], out["opt_g"] = w0, gs
    np.savez_compressed(HERE / "cnn_ops.npz", **out)


def _model_step(ref, x, y, lr, momentum):
    opt = torch.optim.SGD(ref.parameters(), lr=lr, momentum=momentum)
    w0 = {k: v.detach().clone().numpy() for k, v in ref.state_dict().items()}
    logits = ref(x)
    loss = F.cross_entropy(logits, y)
    loss.backward()
    grads = {n: p.grad.clone().numpy() for n, p in ref.named_parameters()}
    opt.step()
    w1 = {n: p.detach().numpy() for n, p in ref.named_parameters()}
    return w0, logits.detach().numpy(), np.float32(loss.item()), grads, w1


def model_steps():
    torch.manual_seed(0)
    g = torch.Generator().manual_seed(1234)
    # LeNet-5 fp32, batch 16 (configs[0] is batch 64; 16 keeps the fixture small)
    ref = R.RefLeNet5()
    x = torch.randn(16, 1, 28, 28, generator=g)
    y = torch.randint(0, 10, (16,), generator=g)
    w0, lg, loss, gr, w1 = _model_step(ref, x, y, 0.05, 0.9)
    out = dict(META, x=x.numpy(), y=y.numpy(), logits=lg, loss=loss)
    out.update({"w0:" + k: v for k, v in w0.items()})
    out.update({"g:" + k: v for k, v in gr.items() if "conv" in k or "fc3" in k})
    out.update({"w1:" + k: v for k, v in w1.items() if "conv" in k or "fc3" in k})
    np.savez_compressed(HERE / "lenet5_step.npz", **out)
    # tiny bottleneck ResNet (base width 16), batch 4, 32x32
    ref = R.RefResNet((1, 1), True, 10, base_width=16)
    x = torch.randn(4, 3, 32, 32, generator=g)
    y = torch.randint(0, 10, (4,), generator=g)
    w0, lg, loss, gr, w1 = _model_step(ref, x, y, 0.1, 0.9)
    out = dict(META, x=x.numpy(), y=y.numpy(), logits=lg, loss=loss)
    out.update({"w0:" + k: v for k, v in w0.items() if "num_batches" not in k})
    out.update({"g:" + k: v for k, v in gr.items()})
    out.update({"bn1_running_var": ref._bn1.running_var.numpy()})
    np.savez_compressed(HERE / "resnet_tiny_step.npz", **out)


def transformer_blocks():
    """Track R: the reference's own classes (fp32, dropout disabled) -- forward, return_kv and gradients."""
    sys.path.insert(0, "/root/reference/src")
    from replicann.arch.transformer import TransformerCrossDecoder, TransformerDecoder, TransformerEncoder
    from replicann.nn.attention import MultiheadCrossAttention, MultiheadSelfAttention
    torch.manual_seed(0)
    g = torch.Generator().manual_seed(1234)
    out = dict(META)
    B, T, E, H = 2, 8, 16, 4
    x = torch.randn(B, T, E, generator=g)

    def pack(tag, mod, outputs, inputs):
        out.update({f"{tag}:sd:{k}": v.detach().numpy() for k, v in mod.state_dict().items()})
        for i, o in enumerate(outputs):
            out[f"{tag}:out{i}"] = o.detach().numpy()
        mod.zero_grad()
        for inp in inputs:
            inp.grad = None
        (outputs[0] * torch.arange(outputs[0].numel(), dtype=torch.float32).reshape(outputs[0].shape).cos()).sum().backward()
        out.update({f"{tag}:grad:{n}": p.grad.numpy() for n, p in mod.named_parameters() if p.grad is not None})
        for i, inp in enumerate(inputs):
            out[f"{tag}:din{i}"] = inp.grad.numpy()

    # NOTE (reference quirk, nn/attention.py:80-83): the multi-head wrappers do not forward p_dropout to their
    # heads, so the per-head attention dropout stays at its default 0.1 whatever the block is given.
    # Deterministic fixtures therefore need .eval() (gradients are still recorded in eval mode).
    enc = TransformerEncoder(H, E, p_dropout=0.0).eval()
    xi = x.clone().requires_grad_()
    pack("enc", enc, [enc(xi)], [xi])
    xi = x.clone().requires_grad_()
    pack("enc_kv", enc, list(enc(xi, return_kv=True)), [xi])
    dec = TransformerDecoder(H, E, context_size=16, p_dropout=0.0).eval()
    xi = x.clone().requires_grad_()
    pack("dec", dec, [dec(xi)], [xi])
    cd = TransformerCrossDecoder(H, E, context_size=16, p_dropout=0.0).eval()
    ek = torch.randn(B, 6, E, generator=g).requires_grad_()
    ev = torch.randn(B, 6, E, generator=g).requires_grad_()
    xi = x.clone().requires_grad_()
    pack("xdec", cd, [cd(xi, ek, ev)], [xi, ek, ev])
    # attention modules with masks: bool (True = masked out), float (additive)
    mha = MultiheadSelfAttention(H, E // H, E, p_dropout=0.0).eval()
    bmask = torch.triu(torch.ones(T, T, dtype=torch.bool), diagonal=1)
    fmask = torch.randn(T, T, generator=g)
    xi = x.clone().requires_grad_()
    pack("mha_bool", mha, [mha(xi, mask=bmask)], [xi])
    xi = x.clone().requires_grad_()
    pack("mha_float", mha, [mha(xi, mask=fmask.clone())], [xi])
    out["mask_bool"], out["mask_float"] = bmask.numpy(), fmask.numpy()
    mca = MultiheadCrossAttention(H, E // H, E, p_dropout=0.0).eval()
    xi = x.clone().requires_grad_()
    pack("mca", mca, [mca(xi, ek.detach(), ev.detach())], [xi])
    # Reference defect (nn/attention.py:164-165): the return_kv branch of MultiheadCrossAttention does not
    # forward return_kv to its heads, so unpacking `z, k_, v_ = h(...)` iterates a [B,T,hs] tensor over the
    # batch dimension: ValueError for every batch size except 3.  Recorded so the drop-in mirrors it.
    try:
        mca(x, ek.detach(), ev.detach(), return_kv=True)
        out["mca_kv_raises"] = np.array("")
    except ValueError as e:
        out["mca_kv_raises"] = np.array(str(e))
    out["x"], out["enc_k"], out["enc_v"] = x.numpy(), ek.detach().numpy(), ev.detach().numpy()
    np.savez_compressed(HERE / "transformer_blocks.npz", **out)


def transformer_stacks():
    """stacks, embeddings and LM loss.  The reference has no such classes; the oracle is a torch-CPU
    network assembled from the reference's OWN block classes (arch/transformer.py:87-201) plus torch.nn.Embedding /
    torch.nn.Linear / F.cross_entropy, with the attribute names of replicann_b200.arch.stacks so that state_dicts
    interchange.  fp32, eval mode (see the dropout note in transformer_blocks)."""
    sys.path.insert(0, "/root/reference/src")
    import torch.nn as nn
    from replicann.arch.transformer import TransformerCrossDecoder, TransformerDecoder, TransformerEncoder
    out = dict(META)

    class Stack(nn.Module):
        def __init__(self, blocks, causal_ctx=None):
            super().__init__()
            self._blocks = nn.ModuleList(blocks)
            if causal_ctx is not None:
                for b in self._blocks:
                    b._attn_mask = torch.triu(torch.ones(causal_ctx, causal_ctx, dtype=torch.bool), diagonal=1)

    class Emb(nn.Module):
        def __init__(self, V, E, ctx):
            super().__init__()
            self._tokens = nn.Embedding(V, E)
            self._positions = nn.Embedding(ctx, E)

        def forward(self, ids):
            return self._tokens(ids) + self._positions(torch.arange(ids.shape[1]))

    class LM(nn.Module):
        def __init__(self, V, L, H, E, ctx, causal, tie):
            super().__init__()
            self._embedding = Emb(V, E, ctx)
            self._decoder = Stack([TransformerDecoder(H, E, context_size=ctx, p_dropout=0.0) for _ in range(L)],
                                  ctx if causal else None)
            self._head = nn.Linear(E, V, bias=not tie)
            if tie:
                self._head.weight = self._embedding._tokens.weight

        def forward(self, ids):
            x = self._embedding(ids)
            for b in self._decoder._blocks:
                x = b(x)
            return self._head(x)

    def record(tag, mod, extra):
        out.update({f"{tag}:sd:{k}": v.detach().numpy() for k, v in mod.state_dict().items()})
        out.update({f"{tag}:grad:{n}": p.grad.numpy() for n, p in mod.named_parameters() if p.grad is not None})
        out.update({f"{tag}:{k}": (v.detach().numpy() if torch.is_tensor(v) else v) for k, v in extra.items()})

    g = torch.Generator().manual_seed(4321)
    V, L, H, E, ctx, B, T = 48, 2, 2, 16, 12, 3, 8
    for tag, causal, tie in (("lm", True, False), ("lm_tied_noncausal", False, True)):
        torch.manual_seed(7)
        lm = LM(V, L, H, E, ctx, causal, tie).eval()
        ids = torch.randint(0, V, (B, T), generator=g)
        ids[0, :3] = 5                       # repeated tokens: several rows scatter into one table row
        tgt = torch.randint(0, V, (B, T), generator=g)
        logits = lm(ids)
        loss = F.cross_entropy(logits.reshape(-1, V), tgt.reshape(-1))
        loss.backward()
        record(tag, lm, dict(ids=ids, targets=tgt, logits=logits, loss=np.float32(loss.item()),
                             cfg=np.array([V, L, H, E, ctx, int(causal), int(tie)])))

    def wsum(o):
        return (o * torch.arange(o.numel(), dtype=torch.float32).reshape(o.shape).cos()).sum()

    torch.manual_seed(8)
    enc = Stack([TransformerEncoder(H, E, p_dropout=0.0) for _ in range(L)]).eval()
    x = torch.randn(B, T, E, generator=g).requires_grad_()
    h = x
    for b in enc._blocks[:-1]:
        h = b(h)
    y, k, v = enc._blocks[-1](h, return_kv=True)
    (wsum(y) + 0.5 * wsum(k) - 0.25 * wsum(v)).backward()
    record("encstack", enc, dict(x=x, y=y, k=k, v=v, dx=x.grad))

    torch.manual_seed(9)
    xd = Stack([TransformerCrossDecoder(H, E, context_size=ctx, p_dropout=0.0) for _ in range(L)], ctx).eval()
    x2 = torch.randn(B, T, E, generator=g).requires_grad_()
    ek = k.detach().clone().requires_grad_()
    ev = v.detach().clone().requires_grad_()
    h = x2
    for b in xd._blocks:
        h = b(h, ek, ev)
    wsum(h).backward()
    record("xdecstack", xd, dict(x=x2, y=h, dx=x2.grad, dek=ek.grad, dev=ev.grad))
    np.savez_compressed(HERE / "transformer_stacks.npz", **out)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "stacks":   # add the round-2 fixture without rewriting the others
        transformer_stacks()
    else:
        cnn_ops()
        model_steps()
        transformer_blocks()
        transformer_stacks()
    for f in sorted(HERE.glob("*.npz")):
        print(f.name, f.stat().st_size)
