"""Model-level GPU parity: a full training step (forward, loss, backward, SGD) of the replicann_b200 CNNs
against independent torch.nn CPU models with identical parameters, inputs and labels.

Tolerances (BASELINE.json north_star): fp32 paths 1e-4 relative; bf16 paths 1e-2 on logits and loss and
gradient cosine similarity >= 0.999."""
import pytest
import torch
import torch.nn.functional as F

import ref_models as R

pytestmark = pytest.mark.gpu


def rel(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return ((a - b).abs().max() / (b.abs().max() + 1e-12)).item()


def cosine(a, b):
    a, b = a.detach().float().cpu().flatten().double(), b.detach().float().cpu().flatten().double()
    return (a @ b / (a.norm() * b.norm() + 1e-300)).item()


def grads_by_name(model):
    return {n: p.grad for n, p in model.named_parameters() if p.grad is not None}


def test_lenet5_fp32_train_step():
    """BASELINE.json configs[0]: LeNet-5 fp32, batch 64, synthetic 1x28x28."""
    from replicann_b200.arch.lenet import LeNet5
    from replicann_b200.optim import SGD
    from replicann_b200.training import Trainer
    torch.manual_seed(0)
    ref = R.RefLeNet5()
    ours = LeNet5().cuda()
    R.load_into(ours, ref)
    g = torch.Generator().manual_seed(1234)
    x = torch.randn(64, 1, 28, 28, generator=g)
    y = torch.randint(0, 10, (64,), generator=g)
    # reference step
    opt_r = torch.optim.SGD(ref.parameters(), lr=0.05, momentum=0.9)
    logits_r = ref(x)
    loss_r = F.cross_entropy(logits_r, y)
    loss_r.backward()
    gr = {n: p.grad.clone() for n, p in ref.named_parameters()}
    opt_r.step()
    # ours
    xc = x.cuda().contiguous(memory_format=torch.channels_last)
    logits_o = ours(xc)
    assert rel(logits_o, logits_r) < 1e-4
    opt = SGD(ours.parameters(), lr=0.05, momentum=0.9)
    tr = Trainer(ours, opt)
    loss_o = tr.step(xc, y.cuda())
    assert abs(loss_o.item() - loss_r.item()) < 1e-4 * max(1, abs(loss_r.item()))
    go = grads_by_name(ours)
    for n, gref in gr.items():
        assert rel(go[n], gref) < 1e-4, n
    for (n, po), (_, pr) in zip(ours.named_parameters(), ref.named_parameters()):
        assert rel(po, pr) < 1e-4, n
    # argmax of the predictions: bit-exact
    assert torch.equal(logits_o.argmax(-1).cpu(), logits_r.argmax(-1))


def _ref_pass(ref, x, y, rounded):
    R.set_rounding(rounded)
    try:
        ref.zero_grad()
        logits = ref(x)
        loss = F.cross_entropy(logits, y)
        loss.backward()
    finally:
        R.set_rounding(False)
    return logits.detach(), loss.item(), {n: p.grad.clone() for n, p in ref.named_parameters()}


def _compare(go, logits_o, loss_o, logits_r, loss_r, gr):
    cos = {n: cosine(go[n], g) for n, g in gr.items() if n in go}
    allo = torch.cat([go[n].float().flatten().cpu() for n in gr])
    allr = torch.cat([g.flatten() for g in gr.values()])
    return {"logits_rel": rel(logits_o, logits_r), "loss_o": loss_o, "loss_r": loss_r,
            "global_cos": cosine(allo, allr), "min_cos": min(cos.values()),
            "low_cos": {n: round(c, 4) for n, c in sorted(cos.items(), key=lambda kv: kv[1])[:4]}}


def _bf16_step_check(ours, ref, x, y, lr=0.05, deep_bn=False, deep_relu=False):
    """GPU bf16 training step vs (a) the fp32 CPU reference and (b) the same reference with activations and
    activation-gradients rounded to bf16 wherever the kernels store bf16 ("bf16-storage reference").

    Networks without BatchNorm (AlexNet, VGG) are held to the north-star tolerance: loss and logits within 1e-2
    of BOTH references; gradient cosine >= 0.999 over the whole gradient and >= 0.99 for every single parameter
    tensor against the bf16-storage reference (same arithmetic, kernels vs torch).  Against the fp32 reference the
    gradient cosine is bounded by what bf16 storage itself costs -- the bf16-storage reference sits at 0.9986
    (AlexNet) from the fp32 one -- so there the assertion is: no farther from fp32 than that reference is
    (within 5e-4), and 0.999 wherever the floor allows it.

    `deep_bn=True` (the ResNets): at random initialisation a deep BatchNorm network amplifies bf16 storage
    rounding chaotically -- the bf16-storage reference itself sits at gradient cosine 0.16-0.96 from the fp32
    reference -- so a whole-network gradient comparison cannot tell a correct kernel from a wrong one.  For
    these networks only the loss is asserted here (it is an average, so it holds 1e-2); logits and gradients are
    asserted, strictly and with no noise-floor allowance, block by block on reference inputs in
    tests/test_resnet_blocks_teacher_forced.py.  The statistics are still printed for the record."""
    from replicann_b200.optim import SGD
    from replicann_b200.training import Trainer
    R.load_into(ours, ref)
    ref.train()
    lr32, loss32, g32 = _ref_pass(ref, x, y, rounded=False)
    for m in ref.modules():  # the second reference pass must not see updated running statistics
        if isinstance(m, torch.nn.BatchNorm2d):
            m.reset_running_stats()
    lrq, lossq, gq = _ref_pass(ref, x, y, rounded=True)
    xc = x.cuda()
    ours.train()
    opt = SGD(ours.parameters(), lr=lr, momentum=0.9)
    logits_o = ours(xc)
    loss_o = Trainer(ours, opt).step(xc, y.cuda()).item()
    go = grads_by_name(ours)
    st_q = _compare(go, logits_o, loss_o, lrq, lossq, gq)
    st_32 = _compare(go, logits_o, loss_o, lr32, loss32, g32)
    floor = _compare(gq, lrq, lossq, lr32, loss32, g32)   # what bf16 STORAGE alone does to the reference
    print("STAT bf16-storage-ref", st_q, flush=True)
    print("STAT fp32-ref", st_32, flush=True)
    print("STAT noise-floor", floor, flush=True)
    assert all(torch.isfinite(g_).all() for g_ in go.values())
    assert abs(st_32["loss_o"] - st_32["loss_r"]) < 1e-2 * max(1.0, abs(st_32["loss_r"])), st_32
    assert abs(st_q["loss_o"] - st_q["loss_r"]) < 1e-2 * max(1.0, abs(st_q["loss_r"])), st_q
    if deep_relu:
        # 16 ReLU layers without normalisation: the two REFERENCES sit at gradient cosine 0.9972 (0.935 for the first
        # layer's filter) from each other -- bf16 storage flips ReLU decisions whose effect grows with depth -- so the
        # assertion is "no farther from either reference than they are from each other"; logits and loss stay strict.
        # Strict per-layer parity at these very shapes: tests/test_tc_real_shapes.py, test_vgg_small_bf16_train_step.
        for st in (st_q, st_32):
            assert st["logits_rel"] < 1e-2, st
            assert st["global_cos"] >= floor["global_cos"] - 2e-3, (st, floor)
            assert st["min_cos"] >= floor["min_cos"] - 3e-2, (st, floor)
    elif not deep_bn:
        for st in (st_q, st_32):
            assert st["logits_rel"] < 1e-2, st
        assert st_q["global_cos"] >= 0.999 and st_q["min_cos"] >= 0.99, st_q
        assert st_32["global_cos"] >= min(0.999, floor["global_cos"] - 5e-4), (st_32, floor)
    return st_q, st_32, floor


def test_resnet_bottleneck_small_bf16():
    from replicann_b200.arch.resnet import ResNet
    torch.manual_seed(0)
    ref = R.RefResNet((1, 1, 1, 1), True, 16)
    ours = ResNet((1, 1, 1, 1), bottleneck=True, n_classes=16).cuda().to(torch.bfloat16)
    g = torch.Generator().manual_seed(1234)
    x = torch.randn(8, 3, 64, 64, generator=g)
    y = torch.randint(0, 16, (8,), generator=g)
    _bf16_step_check(ours, ref, x, y, deep_bn=True)


def test_resnet18_bf16():
    from replicann_b200.arch.resnet import ResNet18
    torch.manual_seed(0)
    ref = R.RefResNet((2, 2, 2, 2), False, 1000)
    ours = ResNet18().cuda().to(torch.bfloat16)
    g = torch.Generator().manual_seed(1234)
    x = torch.randn(8, 3, 96, 96, generator=g)
    y = torch.randint(0, 1000, (8,), generator=g)
    _bf16_step_check(ours, ref, x, y, deep_bn=True)


def test_resnet50_bf16_train_step():
    """The network BASELINE.json's metric is quoted on, at reduced batch (CPU reference time)."""
    from replicann_b200.arch.resnet import ResNet50
    torch.manual_seed(0)
    ref = R.RefResNet((3, 4, 6, 3), True, 1000)
    ours = ResNet50().cuda().to(torch.bfloat16)
    g = torch.Generator().manual_seed(1234)
    x = torch.randn(8, 3, 224, 224, generator=g)
    y = torch.randint(0, 1000, (8,), generator=g)
    _bf16_step_check(ours, ref, x, y, deep_bn=True)


def test_alexnet_bf16_train_step():
    from replicann_b200.arch.alexnet import AlexNet
    torch.manual_seed(0)
    ref = R.RefAlexNet(1000)
    ours = AlexNet(p_dropout=0.0).cuda().to(torch.bfloat16)
    g = torch.Generator().manual_seed(1234)
    x = torch.randn(8, 3, 224, 224, generator=g)
    y = torch.randint(0, 1000, (8,), generator=g)
    _bf16_step_check(ours, ref, x, y)


def test_vgg_small_bf16_train_step():
    from replicann_b200.arch.vgg import VGG
    cfg = (64, "M", 128, "M", 128, 128, "M", 256, "M", 256, "M")
    torch.manual_seed(0)
    ref = R.RefVGG(cfg, 96, hidden=512)
    ours = VGG(cfg, n_classes=96, p_dropout=0.0, hidden_size=512).cuda().to(torch.bfloat16)
    g = torch.Generator().manual_seed(1234)
    x = torch.randn(4, 3, 224, 224, generator=g)
    y = torch.randint(0, 96, (4,), generator=g)
    _bf16_step_check(ours, ref, x, y)


def test_vgg16_bf16_train_step():
    """The real VGG-16 (configuration D, 1000 classes, 4096-wide classifier: BASELINE.json configs[2]) at 224 x 224,
    batch 2: 13 ReLU-fused convolutions with the 1-bit-mask / pooled-resolution ReLU backward, 5 pools, 3 Linear.
    He initialisation on both sides (torch's default shrinks the signal by ~0.6 per layer, which leaves the logits of
    a 16-layer net to the biases)."""
    from replicann_b200.arch.vgg import VGG16
    torch.manual_seed(0)
    ref = R.RefVGG((64, 64, "M", 128, 128, "M", 256, 256, 256, "M", 512, 512, 512, "M", 512, 512, 512, "M"), 1000)
    for m in ref.modules():
        if isinstance(m, (torch.nn.Conv2d, torch.nn.Linear)):
            torch.nn.init.kaiming_normal_(m.weight, nonlinearity="relu")
            torch.nn.init.normal_(m.bias, std=0.05)
    ours = VGG16(p_dropout=0.0).cuda().to(torch.bfloat16)
    g = torch.Generator().manual_seed(1234)
    x = torch.randn(4, 3, 224, 224, generator=g)
    y = torch.randint(0, 1000, (4,), generator=g)
    _bf16_step_check(ours, ref, x, y, lr=1e-3, deep_relu=True)


def test_state_dict_roundtrip_and_eval_mode():
    from replicann_b200.arch.resnet import ResNet
    torch.manual_seed(0)
    ref = R.RefResNet((1, 1), True, 8)
    ours = ResNet((1, 1), bottleneck=True, n_classes=8).cuda().to(torch.bfloat16)
    R.load_into(ours, ref)
    ref.eval(); ours.eval()
    x = torch.randn(4, 3, 64, 64, generator=torch.Generator().manual_seed(5))
    with torch.no_grad():
        assert rel(ours(x.cuda()), ref(x)) < 2e-2
    sd = ours.state_dict()
    assert set(k for k in ref.state_dict()) <= set(sd)


def test_fused_bottleneck_matches_layerwise_path():
    """BottleneckFn (one autograd node, in-place residual-gradient accumulation) vs the layer-by-layer path."""
    from replicann_b200.arch.resnet import ResNet
    torch.manual_seed(0)
    g = torch.Generator().manual_seed(7)
    x = torch.randn(8, 3, 64, 64, generator=g).cuda()
    y = torch.randint(0, 16, (8,), generator=g).cuda()
    res = {}
    for fused in (True, False):
        torch.manual_seed(1)
        m = ResNet((2, 1), bottleneck=True, n_classes=16).cuda().to(torch.bfloat16)
        for mod in m.modules():
            if hasattr(mod, "_fused"):
                mod._fused = fused
        logits = m(x)
        F.cross_entropy(logits.float(), y).backward()
        res[fused] = (logits.detach(), {n: p.grad.clone() for n, p in m.named_parameters()},
                      {k: v.clone() for k, v in m.state_dict().items() if "running" in k})
    # same kernels; only the BatchNorm statistics are summed in a different order (convolution epilogue vs
    # separate pass), so results agree to rounding, not bitwise
    assert rel(res[True][0], res[False][0]) < 1e-2
    for k in res[True][2]:
        assert torch.allclose(res[True][2][k], res[False][2][k], rtol=2e-2, atol=2e-3), k
    for n, gr in res[True][1].items():
        assert cosine(gr, res[False][1][n]) > 0.99, n   # rounding-order noise only (see noise-floor note)


def test_side_stream_weight_gradients_are_bitwise_the_same(monkeypatch):
    """The bottleneck backward issues its weight-gradient kernels on a side stream (joined with a lag of one
    block and at the end of backward).  Same kernels, same inputs: gradients must match the one-stream run
    bit for bit, over several steps so that a missed join (a race) would show."""
    from replicann_b200 import functional as RF
    from replicann_b200.arch.resnet import ResNet
    g = torch.Generator().manual_seed(11)
    x = torch.randn(16, 3, 64, 64, generator=g).cuda()
    y = torch.randint(0, 16, (16,), generator=g).cuda()
    res = {}
    for side in (True, False):
        monkeypatch.setattr(RF, "_side_enabled", side)
        torch.manual_seed(3)
        m = ResNet((2, 2), bottleneck=True, n_classes=16).cuda().to(torch.bfloat16)
        grads = []
        for _ in range(3):
            for p in m.parameters():
                p.grad = None
            F.cross_entropy(m(x).float(), y).backward()
            grads.append({n: p.grad.clone() for n, p in m.named_parameters()})
        torch.cuda.synchronize()
        assert not any(st.pending for st in RF._sides.values())   # everything joined when backward returned
        res[side] = grads
    for a, b in zip(res[True], res[False]):
        for n in a:
            assert torch.equal(a[n], b[n]), n


def test_fused_stem_pool_is_bit_identical_to_the_two_modules():
    """ResNet stem: max_pool(relu(bn(conv))) with the BatchNorm-apply + ReLU folded into the pooling kernel against
    BatchNorm2d(relu=True) + MaxPool2d: same kernels otherwise, so logits, running statistics and every gradient
    must agree bit for bit."""
    from replicann_b200.arch.resnet import ResNet
    g = torch.Generator().manual_seed(21)
    x = torch.randn(8, 3, 96, 96, generator=g).cuda()
    y = torch.randint(0, 16, (8,), generator=g).cuda()
    res = {}
    for fused in (True, False):
        torch.manual_seed(5)
        m = ResNet((1, 1), bottleneck=True, n_classes=16).cuda().to(torch.bfloat16)
        m._fused_stem = fused
        logits = m(x)
        F.cross_entropy(logits.float(), y).backward()
        res[fused] = (logits.detach(), {n: p.grad.clone() for n, p in m.named_parameters()},
                      {k: v.clone() for k, v in m.state_dict().items() if "running" in k or "num_batches" in k})
    assert torch.equal(res[True][0], res[False][0])
    for k in res[True][2]:
        assert torch.equal(res[True][2][k], res[False][2][k]), k
    for n in res[True][1]:
        assert torch.equal(res[True][1][n], res[False][1][n]), n


def test_relu_backward_at_the_pooled_resolution_is_bit_identical():
    """fuse_relu_pool(conv, pool): the convolution's ReLU backward runs inside the pooling layer's backward on the
    pooled-size gradient and the convolution skips its full-size mask pass; fuse_relu_conv(conv, next): it runs in the
    epilogue of the next convolution's dgrad kernel.  AlexNet (overlapping 3x3/2 windows, first
    layer included) and a VGG (2x2/2 windows): logits and every gradient bit for bit against the unfused modules."""
    from replicann_b200.arch.alexnet import AlexNet
    from replicann_b200.arch.vgg import VGG
    from replicann_b200.nn.layers import Conv2d, MaxPool2d
    g = torch.Generator().manual_seed(31)
    cases = [(lambda: AlexNet(n_classes=24, p_dropout=0.0), (4, 3, 224, 224)),
             (lambda: VGG((64, 64, "M", 128, "M", 128, 128, "M"), n_classes=24, p_dropout=0.0, hidden_size=64), (4, 3, 56, 56))]
    for make, shape in cases:
        x = torch.randn(*shape, generator=g).cuda()
        y = torch.randint(0, 24, (shape[0],), generator=g).cuda()
        res = {}
        for fused in (True, False):
            torch.manual_seed(9)
            m = make().cuda().to(torch.bfloat16)
            n_fused = sum(int(c._grad_premasked) + int(c._emit_relu_mask) for c in m.modules() if isinstance(c, Conv2d))
            assert n_fused >= 4
            if not fused:
                for c in m.modules():
                    if isinstance(c, Conv2d):
                        c._grad_premasked = c._emit_relu_mask = False
                    if isinstance(c, MaxPool2d):
                        c._relu_input = False
            logits = m(x)
            F.cross_entropy(logits.float(), y).backward()
            res[fused] = (logits.detach(), {n: p.grad.clone() for n, p in m.named_parameters()})
        assert torch.equal(res[True][0], res[False][0])
        for n in res[True][1]:
            assert torch.equal(res[True][1][n], res[False][1][n]), n
