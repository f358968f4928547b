"""Teacher-forced block-level parity of the ResNets (BASELINE.json configs[3] is ResNet-50).

A whole-network comparison of a deep BatchNorm network at random initialisation measures chaos, not kernels:
bf16 storage rounding of an early activation is amplified block after block until the gradient of the first
layers has little to do with the fp32 one, whichever implementation produced it.  A single block does not
amplify.  So the torch-CPU network (tests/ref_models.py, activations rounded to bf16 at the points where the
kernels store bf16) is run ONCE; for every residual block its input x_i and the gradient dy_i of its output are
recorded; and each GPU block is then fed exactly those tensors, in training mode, through the same code path
the full network uses (BottleneckFn for bottlenecks).  Its output, input gradient and every parameter gradient
are compared with what the reference block produced from the same (x_i, dy_i).  The stem (layout kernel,
packed-pixel first convolution, BatchNorm+ReLU, max-pool) and the head (global average pool, Linear) are
checked the same way.

Tolerances (north_star: "bf16 paths within 1e-2 on logits and loss with gradient cosine similarity at least 0.999"),
on every tensor, with no noise-floor allowance:
  * forward outputs (block outputs, stem output, logits): max error <= 1e-2 of max|ref| AND cosine >= 0.999;
  * gradients (dx and every dW / dgamma / dbeta): cosine >= 0.999 AND relative L2 error <= 5e-2 (the second bound
    catches a wrong scale, which a cosine cannot see).
Why gradients are not held to a max-norm bound: a ReLU whose pre-activation is within one fp32 rounding of zero
can switch between two correct implementations; in a 6.4 M-element tensor a handful do, and each moves ONE
gradient element by the full size of the incoming gradient.  Measured on the first run of this test (round 2):
block-input gradients at cosine 0.99990-0.99999 with single elements off by up to 0.4 of max|ref|.  The printed
`max` column keeps that figure visible.
"""
import pytest
import torch
import torch.nn.functional as F

import ref_models as R

pytestmark = pytest.mark.gpu

REL_TOL = 1e-2      # forward outputs: max error relative to max|ref|
L2_TOL = 5e-2       # gradients: ||got - ref|| / ||ref||
COS_TOL = 0.999


def _rel(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return ((a - b).abs().max() / (b.abs().max() + 1e-30)).item()


def _cos(a, b):
    a, b = a.detach().float().cpu().flatten().double(), b.detach().float().cpu().flatten().double()
    return (a @ b / (a.norm() * b.norm() + 1e-300)).item()


def _record_reference(ref, x, y):
    """One forward + backward of the bf16-storage reference; returns per-block records and stem / head records."""
    blocks = [b for s in ref._stages for b in s]
    rec = {"x": {}, "out": {}, "dx": {}, "dy": {}}

    def fwd_hook(i):
        def hook(_m, inp, out):
            rec["x"][i] = inp[0].detach().clone()
            rec["out"][i] = out.detach().clone()
            inp[0].register_hook(lambda g, i=i: rec["dx"].__setitem__(i, g.detach().clone()))
            out.register_hook(lambda g, i=i: rec["dy"].__setitem__(i, g.detach().clone()))
        return hook

    handles = [b.register_forward_hook(fwd_hook(i)) for i, b in enumerate(blocks)]
    R.set_rounding(True)
    try:
        ref.zero_grad()
        logits = ref(x)
        F.cross_entropy(logits, y).backward()
    finally:
        R.set_rounding(False)
        for h in handles:
            h.remove()
    grads = {n: p.grad.detach().clone() for n, p in ref.named_parameters()}
    return blocks, rec, logits.detach(), grads


def _nets(kind):
    from replicann_b200.arch.resnet import ResNet, ResNet18, ResNet50
    torch.manual_seed(0)
    if kind == "resnet50":
        return R.RefResNet((3, 4, 6, 3), True, 1000), ResNet50(), 224, 1000
    if kind == "resnet18":
        return R.RefResNet((2, 2, 2, 2), False, 1000), ResNet18(), 96, 1000
    return R.RefResNet((1, 1, 1, 1), True, 16), ResNet((1, 1, 1, 1), bottleneck=True, n_classes=16), 64, 16


def _l2(a, b):
    a, b = a.detach().float().cpu().flatten().double(), b.detach().float().cpu().flatten().double()
    return ((a - b).norm() / (b.norm() + 1e-300)).item()


def _check(tag, got, ref, fails, grad=None):
    """grad=None: decided from the tag (" grad" / " dx" are gradients)."""
    if grad is None:
        grad = tag.endswith(" grad") or tag.endswith(" dx")
    r, c, l2 = _rel(got, ref), _cos(got, ref), _l2(got, ref)
    ok = bool(torch.isfinite(got).all()) and c >= COS_TOL and (l2 <= L2_TOL if grad else r <= REL_TOL)
    print(f"STAT {tag}: max {r:.3e} l2 {l2:.3e} cos {c:.6f}{'' if ok else '  <-- FAIL'}", flush=True)
    if not ok:
        fails.append((tag, r, l2, c))


@pytest.mark.parametrize("kind", ["resnet50", "resnet18", "small_bottleneck"])
def test_every_block_matches_reference_given_reference_inputs(kind):
    ref, ours, hw, n_classes = _nets(kind)
    ours = ours.cuda().to(torch.bfloat16)
    R.load_into(ours, ref)
    ref.train()
    ours.train()
    g = torch.Generator().manual_seed(1234)
    x = torch.randn(8, 3, hw, hw, generator=g)
    y = torch.randint(0, n_classes, (8,), generator=g)
    ref_blocks, rec, logits_ref, gref = _record_reference(ref, x, y)
    our_blocks = [b for s in ours._stages for b in s]
    names = [f"_stages.{si}.{bi}" for si, s in enumerate(ours._stages) for bi in range(len(s))]
    assert len(our_blocks) == len(ref_blocks) == len(rec["dy"])
    fails: list = []

    def to_dev(t):
        return t.to(torch.bfloat16).cuda().contiguous(memory_format=torch.channels_last)

    # ---- residual blocks
    for i, (blk, name) in enumerate(zip(our_blocks, names)):
        for p in blk.parameters():
            p.grad = None
        xi = to_dev(rec["x"][i]).requires_grad_(True)
        out = blk(xi)
        out.backward(to_dev(rec["dy"][i]))
        torch.cuda.synchronize()
        _check(f"{kind} {name} out", out, rec["out"][i], fails)
        _check(f"{kind} {name} dx", xi.grad, rec["dx"][i], fails)
        for pn, p in blk.named_parameters():
            _check(f"{kind} {name}.{pn} grad", p.grad, gref[f"{name}.{pn}"], fails)

    # ---- stem: images -> conv1 -> BatchNorm + ReLU -> max-pool, given d(pool output) = dx of the first block
    from replicann_b200.arch._stem import prepare_input
    for p in list(ours._conv1.parameters()) + list(ours._bn1.parameters()):
        p.grad = None
    z = ours._maxpool(ours._bn1(ours._conv1(prepare_input(x.cuda(), torch.bfloat16))))
    z.backward(to_dev(rec["dx"][0]))
    torch.cuda.synchronize()
    _check(f"{kind} stem out", z, rec["x"][0], fails)
    for pn in ("_conv1.weight", "_bn1.weight", "_bn1.bias"):
        _check(f"{kind} stem {pn} grad", ours.get_parameter(pn).grad, gref[pn], fails)

    # ---- head: global average pool -> Linear, given the last block's output and dlogits of the reference
    for p in ours._fc.parameters():
        p.grad = None
    last = len(our_blocks) - 1
    xl = to_dev(rec["out"][last]).requires_grad_(True)
    pooled = ours._avgpool(xl)
    logits = ours._fc(pooled.reshape(pooled.shape[0], -1))
    dlogits = (torch.softmax(logits_ref, -1) - F.one_hot(y, n_classes).float()) / y.numel()
    logits.backward(dlogits.to(torch.bfloat16).cuda())
    torch.cuda.synchronize()
    _check(f"{kind} head logits", logits, logits_ref, fails)
    _check(f"{kind} head dx", xl.grad, rec["dy"][last], fails)
    _check(f"{kind} head _fc.weight grad", ours._fc.weight.grad, gref["_fc.weight"], fails)
    _check(f"{kind} head _fc.bias grad", ours._fc.bias.grad, gref["_fc.bias"], fails)
    assert not fails, fails
