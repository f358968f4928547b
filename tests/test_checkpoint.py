"""Checkpoint interchange: `.pt` files written by either side load into the other.

  * a genuine reference checkpoint (tests/golden/ref_blocks.pt, written by the reference's own classes through
    torch.save, see tests/golden/make_checkpoint_fixture.py) loads into the drop-in blocks with strict=True,
    including the persisted `_attn_mask` buffer (arch/transformer.py:142-145,176-179);
  * where the reference sources are present (/root/reference, this container only) the full circle is run:
    ours -> .pt -> reference class (strict) -> .pt -> ours, bit for bit;
  * the CNNs interchange with plain torch.nn modules the same way;
  * optimizer state: the flat fp32 buckets convert to and from torch.optim's per-parameter checkpoint format.

Host-side only (state_dict bookkeeping, no kernels), so this runs in the CPU suite.
"""
import sys
from pathlib import Path

import pytest
import torch

G = Path(__file__).resolve().parent / "golden"
REF_SRC = Path("/root/reference/src")


def _ours(kind, cfg, **kw):
    from replicann_b200.arch.transformer import TransformerCrossDecoder, TransformerDecoder, TransformerEncoder
    H, E, ctx = cfg["n_heads"], cfg["embedding_size"], cfg["context_size"]
    if kind == "encoder":
        return TransformerEncoder(H, E, **kw)
    if kind == "decoder":
        return TransformerDecoder(H, E, context_size=ctx, **kw)
    return TransformerCrossDecoder(H, E, context_size=ctx, head_bias=True, **kw)


@pytest.mark.parametrize("kind", ["encoder", "decoder", "cross_decoder"])
def test_reference_pt_loads_strictly(kind):
    ck = torch.load(G / "ref_blocks.pt")
    ours = _ours(kind, ck["config"])
    res = ours.load_state_dict(ck[kind], strict=True)
    assert not res.missing_keys and not res.unexpected_keys
    sd = ours.state_dict()
    assert list(sd) == list(ck[kind])                      # same keys in the same order
    for k, v in ck[kind].items():
        assert sd[k].dtype == v.dtype and torch.equal(sd[k], v), k
    if kind != "encoder":
        ctx = ck["config"]["context_size"]
        assert torch.equal(sd["_attn_mask"], torch.tril(torch.ones(ctx, ctx)))


@pytest.mark.skipif(not REF_SRC.exists(), reason="reference sources are only present in the build container")
@pytest.mark.parametrize("kind", ["encoder", "decoder", "cross_decoder"])
def test_pt_round_trip_through_the_reference_classes(kind, tmp_path):
    sys.path.insert(0, str(REF_SRC))
    try:
        from replicann.arch import transformer as RT
    finally:
        sys.path.remove(str(REF_SRC))
    cfg = {"n_heads": 2, "embedding_size": 16, "context_size": 12}
    torch.manual_seed(3)
    ours = _ours(kind, cfg)
    torch.save(ours.state_dict(), tmp_path / "ours.pt")
    H, E, ctx = cfg["n_heads"], cfg["embedding_size"], cfg["context_size"]
    ref = {"encoder": lambda: RT.TransformerEncoder(H, E),
           "decoder": lambda: RT.TransformerDecoder(H, E, context_size=ctx),
           "cross_decoder": lambda: RT.TransformerCrossDecoder(H, E, context_size=ctx, head_bias=True)}[kind]()
    ref.load_state_dict(torch.load(tmp_path / "ours.pt"), strict=True)
    torch.save(ref.state_dict(), tmp_path / "ref.pt")
    back = _ours(kind, cfg)
    back.load_state_dict(torch.load(tmp_path / "ref.pt"), strict=True)
    for (k, a), (k2, b) in zip(ours.state_dict().items(), back.state_dict().items()):
        assert k == k2 and torch.equal(a, b), k


def test_cnn_pt_round_trip_with_torch_nn(tmp_path):
    import ref_models as R
    from replicann_b200.arch.resnet import ResNet
    torch.manual_seed(4)
    ours = ResNet((1, 1), bottleneck=True, n_classes=8)
    ref = R.RefResNet((1, 1), True, 8)
    torch.save(ours.state_dict(), tmp_path / "ours.pt")
    ref.load_state_dict(torch.load(tmp_path / "ours.pt"), strict=True)
    torch.save(ref.state_dict(), tmp_path / "ref.pt")
    back = ResNet((1, 1), bottleneck=True, n_classes=8)
    back.load_state_dict(torch.load(tmp_path / "ref.pt"), strict=True)
    for (k, a), (k2, b) in zip(ours.state_dict().items(), back.state_dict().items()):
        assert k == k2 and torch.equal(a, b), k
    # a filter that came through torch.nn (NCHW-contiguous) is KRSC again, the layout the kernels read in place
    assert back._conv1.weight.permute(0, 2, 3, 1).is_contiguous()


@pytest.mark.parametrize("which", ["sgd", "adamw"])
def test_optimizer_state_converts_to_and_from_torch_optim(which, tmp_path):
    """Flat-bucket state <-> per-parameter views, through a .pt file and a real torch.optim optimizer."""
    from replicann_b200.arch.transformer import TransformerDecoder
    from replicann_b200.optim import SGD, AdamW
    torch.manual_seed(5)
    ours = TransformerDecoder(2, 16, context_size=8)
    twin = TransformerDecoder(2, 16, context_size=8)
    twin.load_state_dict(ours.state_dict())
    if which == "sgd":
        opt = SGD(ours.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-4)
        names = ("momentum_buffer",)
        t_opt = torch.optim.SGD(twin.parameters(), lr=1.0, momentum=0.5)
    else:
        opt = AdamW(ours.parameters(), lr=2e-3, betas=(0.8, 0.95), weight_decay=0.02)
        names = ("exp_avg", "exp_avg_sq")
        t_opt = torch.optim.AdamW(twin.parameters())
    for g in opt.groups:
        for n in names:
            g.zeros_state(n).uniform_(0.1, 1.0)
    opt._steps = 5
    torch.save(opt.torch_state_dict(), tmp_path / "opt.pt")
    t_opt.load_state_dict(torch.load(tmp_path / "opt.pt"))
    assert t_opt.param_groups[0]["lr"] == opt.lr
    plist = [p for p in ours.parameters() if p.requires_grad]
    for p_ours, p_twin in zip(plist, twin.parameters()):
        g, o = p_ours._rnn_group, p_ours._rnn_offset
        for n in names:
            assert torch.equal(t_opt.state[p_twin][n], g._view(g.state[n], p_ours, o)), n
        if which == "adamw":
            assert float(t_opt.state[p_twin]["step"]) == 5.0
    # and back into a fresh flat optimizer
    torch.save(t_opt.state_dict(), tmp_path / "topt.pt")
    other = TransformerDecoder(2, 16, context_size=8)
    other.load_state_dict(ours.state_dict())
    opt2 = (SGD(other.parameters(), lr=0.0) if which == "sgd" else AdamW(other.parameters()))
    opt2.load_torch_state_dict(torch.load(tmp_path / "topt.pt"))
    assert opt2.lr == opt.lr and opt2._steps >= 1
    for pa, pb in zip(ours.parameters(), other.parameters()):   # per parameter: alignment padding is not state
        ga, gb = pa._rnn_group, pb._rnn_group
        for n in names:
            assert torch.equal(ga._view(ga.state[n], pa, pa._rnn_offset), gb._view(gb.state[n], pb, pb._rnn_offset)), n
    for g2 in opt2.groups:
        assert torch.equal(g2.master, g2.work.float())


def test_packed_projection_weights_are_views_of_the_flat_buffer():
    """The per-head q/k/v Linear weights of a multi-head module are laid out back to back by the flat optimizer
    (optim.ParamPack), so one view of the flat buffer is the packed [3*H*hs, E] projection matrix and one view of
    the flat gradient buffer receives its gradient -- while state_dict keys stay the reference's per-head ones."""
    from replicann_b200.nn.attention import MultiheadCrossAttention, MultiheadSelfAttention
    from replicann_b200.optim import AdamW
    torch.manual_seed(6)
    sa = MultiheadSelfAttention(4, 16, 64, head_bias=True)
    ca = MultiheadCrossAttention(4, 16, 64)
    assert sa._wpack.views() is None                      # no optimizer yet: the concatenating path is used
    AdamW(list(sa.parameters()) + list(ca.parameters()))
    w, gw = sa._wpack.views()
    b, gb = sa._bpack.views()
    heads = sa._heads
    expect = torch.cat([getattr(h, n).weight for n in ("_query", "_key", "_value") for h in heads])
    assert tuple(w.shape) == (3 * 4 * 16, 64) and torch.equal(w, expect)
    assert torch.equal(b, torch.cat([getattr(h, n).bias for n in ("_query", "_key", "_value") for h in heads]))
    gw.fill_(3.0)
    assert all(float(getattr(h, n).weight.grad.min()) == 3.0 for n in ("_query", "_key", "_value") for h in heads)
    wq, _ = ca._wpack.views()
    assert torch.equal(wq, torch.cat([h._query.weight for h in ca._heads]))
    assert "_heads.0._query.weight" in sa.state_dict()
    # zero_grad really clears (gradients arriving through stock autograd accumulate into these views) ...
    opt = AdamW(sa.parameters())
    sa._heads[0]._query.weight.grad.fill_(7.0)
    opt.zero_grad()
    assert float(sa._heads[0]._query.weight.grad.abs().max()) == 0.0
    # ... and a .grad detached by model.zero_grad(set_to_none=True) is re-attached to the flat buffer
    sa.zero_grad(set_to_none=True)
    assert sa._heads[0]._query.weight.grad is None
    opt.zero_grad()
    g0 = sa._heads[0]._query.weight
    assert g0.grad is not None and g0.grad.data_ptr() == g0._rnn_group.grad.data_ptr() + g0._rnn_offset * 4
