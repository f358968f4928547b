"""GPU parity of the HBM-bound kernels and the fp32 path against torch CPU (the dependency the reference
calls into), through the C ABI.  Tolerances: fp32 1e-4 relative, bf16 1e-2 relative (north_star);
indices bit-exact."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def RF():
    from replicann_b200 import functional
    return functional


def rel_err(got, ref):
    got = got.detach().float().cpu()
    ref = ref.detach().float().cpu()
    return ((got - ref).abs().max() / (ref.abs().max() + 1e-12)).item()


def rnd(shape, seed, dtype=torch.float32, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(shape, generator=g) * scale).to(dtype)


# ------------------------------------------------------------------------------------------ BatchNorm
@pytest.mark.parametrize("M,C", [(2 * 14 * 14, 64), (3 * 7 * 7, 256), (1000, 2048), (513, 192), (4096, 8),
                                 (8 * 56 * 56, 64), (40000, 256)])
@pytest.mark.parametrize("relu,res", [(True, False), (True, True), (False, False)])
def test_batchnorm_fwd_bwd(M, C, relu, res):
    f = RF()
    x = rnd((M, C), 1, torch.bfloat16) * 2 + 0.5
    gamma = torch.rand(C, generator=torch.Generator().manual_seed(2)) + 0.5
    beta = rnd((C,), 3)
    resid = rnd((M, C), 4, torch.bfloat16) if res else None
    dy = rnd((M, C), 5, torch.bfloat16)
    rm, rv = torch.zeros(C), torch.ones(C)
    xd = x.cuda()
    rmd, rvd = rm.cuda(), rv.cuda()
    gd, bd = gamma.cuda(), beta.cuda()
    mean, invstd, scale, shift = f.bn_stats(xd, gd, bd, 1e-5, 0.1, rmd, rvd)
    y, bits = f.bn_apply(xd, scale, shift, resid.cuda() if res else None, relu, want_mask=True)
    if relu:  # the 1-bit mask is exactly (stored output > 0)
        exp_bits = (y > 0).reshape(M, C // 8, 8).to(torch.int32)
        weights = (2 ** torch.arange(8, device="cuda", dtype=torch.int32)).reshape(1, 1, 8)
        assert torch.equal(bits.to(torch.int32), (exp_bits * weights).sum(-1))

    # reference: torch BatchNorm on [M, C, 1, 1]
    xr = x.float().reshape(M, C, 1, 1).requires_grad_(True)
    gr = gamma.clone().requires_grad_(True)
    br = beta.clone().requires_grad_(True)
    rr = resid.float().reshape(M, C, 1, 1).requires_grad_(True) if res else None
    rm_ref, rv_ref = rm.clone(), rv.clone()
    out = F.batch_norm(xr, rm_ref, rv_ref, gr, br, training=True, momentum=0.1, eps=1e-5)
    if res:
        out = out + rr
    if relu:
        # A pre-activation within fp32 rounding of zero may land on either side of the ReLU; such ties
        # must be rare, and the backward comparison uses the GPU's own mask so it is tie-free.
        mask = (y.float().cpu() > 0).reshape(M, C, 1, 1)
        assert ((out > 0) != mask).float().mean().item() < 1e-5
        out = out * mask
    out.backward(dy.float().reshape(M, C, 1, 1))
    assert rel_err(y, out.reshape(M, C)) < 1e-2
    assert rel_err(rmd, rm_ref) < 1e-4 and rel_err(rvd, rv_ref) < 1e-3
    dgamma = torch.empty(C, device="cuda")
    dbeta = torch.empty(C, device="cuda")
    # ReLU mask: from the saved output when a residual was added, else recomputed from x (scale, shift)
    from_x = relu and not res
    dx, dres = f.bn_bwd(dy.cuda(), xd, y if (relu and not from_x) else None, mean, invstd, gd, want_dres=res,
                        dgamma=dgamma, dbeta=dbeta, accumulate=False, fwd_scale=scale if from_x else None,
                        fwd_shift=shift if from_x else None)
    if relu and not from_x:  # same backward driven by the bit mask instead of the saved output
        dg2, db2 = torch.empty(C, device="cuda"), torch.empty(C, device="cuda")
        dx2, dres2 = f.bn_bwd(dy.cuda(), xd, None, mean, invstd, gd, want_dres=res, dgamma=dg2, dbeta=db2,
                              accumulate=False, relu_mask=bits)
        assert torch.equal(dx2, dx) and torch.equal(dres2, dres) and torch.equal(dg2, dgamma)
    assert rel_err(dx, xr.grad.reshape(M, C)) < 1e-2
    assert rel_err(dgamma, gr.grad) < 1e-2 and rel_err(dbeta, br.grad) < 1e-2
    if res:
        assert rel_err(dres, rr.grad.reshape(M, C)) < 1e-2


def test_batchnorm_eval_prepare():
    f = RF()
    C = 64
    g, b, rm = rnd((C,), 1), rnd((C,), 2), rnd((C,), 3)
    rv = torch.rand(C) + 0.1
    scale, shift = f.bn_eval_prepare(g.cuda(), b.cuda(), rm.cuda(), rv.cuda(), 1e-5)
    x = rnd((10, C), 4)
    ref = F.batch_norm(x.reshape(10, C, 1, 1), rm, rv, g, b, training=False, eps=1e-5).reshape(10, C)
    assert rel_err(x * scale.cpu() + shift.cpu(), ref) < 1e-5


# ------------------------------------------------------------------------------------------ LayerNorm
@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-4), (torch.bfloat16, 1e-2)])
@pytest.mark.parametrize("M,E,res", [(64, 512, True), (37, 768, False), (8, 16, True)])
def test_layernorm(dtype, tol, M, E, res):
    f = RF()
    x = rnd((M, E), 1, dtype)
    r = rnd((M, E), 2, dtype) if res else None
    g = (torch.rand(E) + 0.5).to(dtype)
    b = rnd((E,), 3, dtype)
    dy = rnd((M, E), 4, dtype)
    xr = x.float().requires_grad_(True)
    rr = r.float().requires_grad_(True) if res else None
    gr, br = g.float().requires_grad_(True), b.float().requires_grad_(True)
    s = xr + rr if res else xr
    ref = F.layer_norm(s, (E,), gr, br, 1e-5)
    ref.backward(dy.float())
    y, xin, mean, rstd = f.layernorm_fwd(x.cuda(), r.cuda() if res else None, g.cuda(), b.cuda(), 1e-5)
    assert rel_err(y, ref) < tol
    dg = torch.empty(E, dtype=dtype, device="cuda")
    db = torch.empty(E, dtype=dtype, device="cuda")
    dx = f.layernorm_bwd(dy.cuda(), xin, g.cuda(), mean, rstd, dg, db, False)
    assert rel_err(dx, xr.grad) < tol
    assert rel_err(dg, gr.grad) < max(tol, 2e-4) and rel_err(db, br.grad) < max(tol, 2e-4)


# ------------------------------------------------------------------------------------------ pooling
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("N,H,W,C,k,s,p", [(2, 112, 112, 64, 3, 2, 1), (2, 13, 13, 16, 3, 2, 0), (3, 28, 28, 8, 2, 2, 0),
                                            (1, 7, 9, 24, 3, 1, 1), (2, 9, 7, 16, 2, 2, 0), (2, 15, 11, 8, 3, 2, 1)])
def test_maxpool_indices_bit_exact(dtype, N, H, W, C, k, s, p):
    f = RF()
    x = rnd((N, H, W, C), 1, dtype)
    # ties on purpose: quantise so that equal maxima occur inside windows
    x = (x * 2).round() / 2
    xr = x.float().permute(0, 3, 1, 2).contiguous().requires_grad_(True)
    ref, ridx = F.max_pool2d(xr, k, s, p, return_indices=True)
    y, idx, widx = f.maxpool2d_fwd(x.cuda(), k, s, p, return_window_idx=True)
    assert torch.equal(y.float().cpu(), ref.permute(0, 2, 3, 1))
    assert torch.equal(idx.cpu().long(), ridx.permute(0, 2, 3, 1))  # bit-exact argmax
    dy = rnd(tuple(y.shape), 2, dtype)
    ref.backward(dy.float().permute(0, 3, 1, 2))
    tol = 1e-5 if dtype == torch.float32 else 1e-2
    dx = f.maxpool2d_bwd(dy.cuda(), idx, tuple(x.shape), k, s, p)
    assert rel_err(dx, xr.grad.permute(0, 2, 3, 1)) < tol
    dx2 = f.maxpool2d_bwd(dy.cuda(), None, tuple(x.shape), k, s, p, window_idx=widx)  # 1-byte argmax form
    assert torch.equal(dx2, dx)


@pytest.mark.parametrize("pad", [1, 0])
@pytest.mark.parametrize("N,H,W,C", [(2, 30, 18, 64), (3, 13, 11, 16), (1, 8, 8, 8), (2, 21, 112, 64), (2, 55, 55, 64),
                                     (2, 14, 27, 192)])
def test_maxpool_3x3s2_marching_kernels_match_the_row_kernels(N, H, W, C, pad, monkeypatch):
    """The marching forward / backward kernels of the 3x3 / stride 2 / pad 1 pool (rows carried in registers) against
    the one-row-per-CTA kernels: values, both index forms and dx bit for bit -- with ties, -inf, NaN, and the
    BatchNorm+ReLU-fused forward; and against torch for the NaN-free part."""
    f = RF()
    x = (rnd((N, H, W, C), 5, torch.bfloat16) * 2).round() / 2
    x[0, 2, 3, 1] = float("nan"); x[0, 3, 3, 1] = float("nan"); x[-1, H - 1, W - 1, 0] = float("nan")
    x[0, :5, :5, 2] = float("-inf")
    scale = (torch.rand(C) + 0.5).cuda(); shift = (torch.randn(C) * 0.3).cuda()
    dy = rnd((N, (H + 2 * pad - 3) // 2 + 1, (W + 2 * pad - 3) // 2 + 1, C), 6, torch.bfloat16).cuda()
    out = {}
    for mode in ("0", "1"):
        monkeypatch.setenv("RNN_POOL_MARCH", mode)
        y, idx, widx = f.maxpool2d_fwd(x.cuda(), 3, 2, pad, return_window_idx=True)
        yb, wb = f.bn_relu_maxpool2d_fwd(x.cuda(), scale, shift, 3, 2, pad)
        dx = f.maxpool2d_bwd(dy, None, tuple(x.shape), 3, 2, pad, window_idx=widx)
        out[mode] = (y, idx, widx, yb, wb, dx)
    for name, a, b in zip(("y", "idx", "widx", "y_bn", "widx_bn", "dx"), out["0"], out["1"]):
        if a.dtype == torch.bfloat16:   # bit for bit, except that a NaN may carry a different payload
            assert torch.equal(torch.isnan(a), torch.isnan(b)), name
            a, b = torch.nan_to_num(a.float(), nan=0.0), torch.nan_to_num(b.float(), nan=0.0)
            a, b = a.view(torch.int32), b.view(torch.int32)
        assert torch.equal(a, b), name
    ref, ridx = F.max_pool2d(x.float().permute(0, 3, 1, 2), 3, 2, pad, return_indices=True)
    assert torch.equal(out["1"][1].cpu().long(), ridx.permute(0, 2, 3, 1))


@pytest.mark.parametrize("march", ["1", "0"])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("N,H,W,C,k,s,p", [(2, 28, 28, 64, 2, 2, 0), (2, 27, 27, 64, 3, 2, 0), (2, 30, 18, 16, 3, 2, 1),
                                            (1, 9, 7, 12, 2, 2, 0), (2, 12, 12, 6, 3, 1, 1)])
def test_maxpool_backward_with_the_relu_mask_of_the_pooled_output(dtype, N, H, W, C, k, s, p, march, monkeypatch):
    """rnn_maxpool2d_relu_bwd: max_pool2d backward with the ReLU in front of the pool folded in (mask = pooled y > 0)
    against relu_bwd at the pooled size followed by the plain kernel -- bit for bit, every kernel variant."""
    monkeypatch.setenv("RNN_POOL_MARCH", march)
    f = RF()
    x = torch.relu(rnd((N, H, W, C), 7, dtype)).cuda()           # a ReLU output: many exact zeros
    y, _, widx = f.maxpool2d_fwd(x, k, s, p, return_indices=False, return_window_idx=True)
    dy = rnd(tuple(y.shape), 8, dtype).cuda()
    ref = f.maxpool2d_bwd(f.relu_bwd(dy, y), None, tuple(x.shape), k, s, p, window_idx=widx)
    got = f.maxpool2d_bwd(dy, None, tuple(x.shape), k, s, p, window_idx=widx, relu_y=y)
    assert torch.equal(got, ref)
    # and it is the gradient of max_pool2d(relu(.)) at a pre-activation whose ReLU is x
    pre = x.float().cpu().clone()
    pre[pre == 0] = -1.0
    pre = pre.permute(0, 3, 1, 2).requires_grad_(True)
    F.max_pool2d(torch.relu(pre), k, s, p).backward(dy.float().cpu().permute(0, 3, 1, 2))
    assert rel_err(got, pre.grad.permute(0, 2, 3, 1)) < (1e-6 if dtype == torch.float32 else 1e-2)


def test_maxpool_nan_propagates():
    f = RF()
    x = rnd((1, 4, 4, 8), 1)
    x[0, 1, 1, 3] = float("nan")
    ref, ridx = F.max_pool2d(x.permute(0, 3, 1, 2), 2, 2, 0, return_indices=True)
    y, idx, _ = f.maxpool2d_fwd(x.cuda(), 2, 2, 0)
    assert torch.equal(idx.cpu().long(), ridx.permute(0, 2, 3, 1))
    assert torch.isnan(y.cpu()[0, 0, 0, 3])


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-5), (torch.bfloat16, 1e-2)])
@pytest.mark.parametrize("N,H,W,C,k,s,p", [(2, 28, 28, 8, 2, 2, 0), (2, 10, 10, 16, 2, 2, 0), (1, 9, 9, 8, 3, 2, 1)])
def test_avgpool(dtype, tol, N, H, W, C, k, s, p):
    f = RF()
    x = rnd((N, H, W, C), 1, dtype)
    xr = x.float().permute(0, 3, 1, 2).contiguous().requires_grad_(True)
    ref = F.avg_pool2d(xr, k, s, p)
    y = f.avgpool2d_fwd(x.cuda(), k, s, p)
    assert rel_err(y, ref.permute(0, 2, 3, 1)) < tol
    dy = rnd(tuple(y.shape), 2, dtype)
    ref.backward(dy.float().permute(0, 3, 1, 2))
    dx = f.avgpool2d_bwd(dy.cuda(), tuple(x.shape), k, s, p)
    assert rel_err(dx, xr.grad.permute(0, 2, 3, 1)) < tol


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-5), (torch.bfloat16, 1e-2)])
@pytest.mark.parametrize("N,HW,C", [(4, 49, 2048), (3, 36, 256), (2, 5, 16)])
def test_global_avgpool(dtype, tol, N, HW, C):
    f = RF()
    x = rnd((N, HW, 1, C), 1, dtype)
    ref = x.float().mean(dim=(1, 2))
    y = f.global_avgpool_fwd(x.cuda())
    assert rel_err(y.reshape(N, C), ref) < tol
    dy = rnd((N, 1, 1, C), 2, dtype)
    dx = f.global_avgpool_bwd(dy.cuda(), tuple(x.shape))
    assert rel_err(dx, (dy.float() / HW).expand(N, HW, 1, C)) < tol


# ------------------------------------------------------------------------------------------ softmax / CE
@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-5), (torch.bfloat16, 1e-2)])
@pytest.mark.parametrize("rows,cols", [(64, 128), (7, 1000), (5, 10)])
def test_softmax(dtype, tol, rows, cols):
    f = RF()
    x = rnd((rows, cols), 1, dtype, 3.0)
    xr = x.float().requires_grad_(True)
    ref = F.softmax(xr, dim=-1)
    y = f.softmax_fwd(x.cuda())
    assert rel_err(y, ref) < tol
    dy = rnd((rows, cols), 2, dtype)
    ref.backward(dy.float())
    yq = y if dtype == torch.float32 else ref.detach().to(dtype).cuda()
    dx = f.softmax_bwd(dy.cuda(), yq)
    assert rel_err(dx, xr.grad) < max(tol, 1e-4)


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-5), (torch.bfloat16, 1e-2)])
@pytest.mark.parametrize("rows,cols", [(256, 1000), (64, 10), (3, 7)])
def test_softmax_cross_entropy(dtype, tol, rows, cols):
    f = RF()
    x = rnd((rows, cols), 1, dtype, 2.0)
    t = torch.randint(0, cols, (rows,), generator=torch.Generator().manual_seed(2))
    xr = x.float().requires_grad_(True)
    ref = F.cross_entropy(xr, t)
    ref.backward()
    loss, dl = f.softmax_ce(x.cuda(), t.cuda())
    assert abs(loss.item() - ref.item()) < 1e-4 * max(1.0, abs(ref.item()))
    assert rel_err(dl, xr.grad) < tol
    am = f.argmax(x.cuda())
    assert torch.equal(am.cpu(), x.float().argmax(dim=-1))  # bit-exact


# ------------------------------------------------------------------------------------------ optimizers
@pytest.mark.parametrize("gdt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("nesterov", [False, True])
def test_sgd_matches_torch(gdt, nesterov):
    f = RF()
    n = 4096 + 64
    w0 = rnd((n,), 1)
    p = torch.nn.Parameter(w0.clone())
    opt = torch.optim.SGD([p], lr=0.1, momentum=0.9, weight_decay=1e-4, nesterov=nesterov)
    w = w0.clone().cuda()
    buf = torch.zeros(n, device="cuda")
    wc = torch.empty(n, dtype=torch.bfloat16, device="cuda")
    for step in range(3):
        g = rnd((n,), 10 + step).to(gdt)
        p.grad = g.float().clone()
        opt.step()
        f.sgd_step(w, buf, g.cuda(), wc, lr=0.1, momentum=0.9, weight_decay=1e-4, nesterov=nesterov,
                   first_step=(step == 0))
    assert rel_err(w, p.data) < 1e-5
    assert rel_err(wc, p.data) < 1e-2


def test_adamw_matches_torch():
    f = RF()
    n = 2048
    w0 = rnd((n,), 1)
    p = torch.nn.Parameter(w0.clone())
    opt = torch.optim.AdamW([p], lr=1e-2, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2)
    w = w0.clone().cuda()
    m = torch.zeros(n, device="cuda")
    v = torch.zeros(n, device="cuda")
    for step in range(1, 4):
        g = rnd((n,), 10 + step)
        p.grad = g.clone()
        opt.step()
        f.adamw_step(w, m, v, g.cuda(), None, lr=1e-2, weight_decay=1e-2, step=step)
    assert rel_err(w, p.data) < 1e-5


# ------------------------------------------------------------------------------------------ elementwise
def test_layout_roundtrip_and_padding():
    f = RF()
    x = rnd((2, 3, 17, 20), 1)
    y = f.nchw_to_nhwc_bf16(x.cuda(), cpad=8)
    assert tuple(y.shape) == (2, 17, 20, 8)
    assert torch.equal(y[..., :3].float().cpu(), x.to(torch.bfloat16).float().permute(0, 2, 3, 1))
    assert (y[..., 3:] == 0).all()
    back = f.nhwc_bf16_to_nchw_f32(y, channels=3)
    assert torch.equal(back.cpu(), x.to(torch.bfloat16).float())


@pytest.mark.parametrize("N,C,H,W,R,S,stride,pad", [
    (4, 3, 224, 224, 7, 7, 2, 3),     # ResNet stem: the image fast path (W % 4 == 0, 8 padded rows per block)
    (2, 3, 67, 67, 11, 11, 4, 2),     # W % 4 != 0: generic padded kernel
    (3, 1, 28, 28, 5, 5, 1, 2),       # one channel
    (2, 3, 30, 36, 3, 3, 1, 1),       # Hp not a multiple of the rows per block
])
@pytest.mark.parametrize("norm", [False, True])
def test_input_layout_kernel_with_normalisation(N, C, H, W, R, S, stride, pad, norm):
    """fp32 NCHW -> bf16 NHWC (+ per-channel normalisation), into the zero-padded packed-pixel buffer
    of the first convolution and into the plain channel-padded layout.  Bit-exact against torch: same fp32
    arithmetic ((x - mean) * (1/std)), same round-to-nearest-even to bf16, exact zeros in every padding."""
    from replicann_b200._lib import conv_desc
    f = RF()
    x = rnd((N, C, H, W), 5, scale=0.5) + 0.4
    mean = torch.tensor([0.485, 0.456, 0.406][:C])
    inv_std = 1.0 / torch.tensor([0.229, 0.224, 0.225][:C])
    ref = x
    kw = {}
    if norm:
        ref = (x - mean.view(1, C, 1, 1)) * inv_std.view(1, C, 1, 1)
        kw = dict(mean=mean.cuda(), inv_std=inv_std.cuda())
    ref = ref.to(torch.bfloat16).permute(0, 2, 3, 1)
    d = conv_desc(N, H, W, C, 64, R, S, stride, pad)
    hp, wp, _ = f.first_conv_geometry(d)
    y = f.pack_first_input(x.cuda(), d, **kw).cpu()
    assert tuple(y.shape) == (N, hp, wp, 8)
    assert torch.equal(y[:, pad:pad + H, pad:pad + W, :C], ref)
    border = y.clone()
    border[:, pad:pad + H, pad:pad + W, :C] = 0
    assert (border == 0).all()
    y2 = f.nchw_to_nhwc_bf16(x.cuda(), cpad=8, **kw).cpu()
    assert torch.equal(y2[..., :C], ref) and (y2[..., C:] == 0).all()


def test_first_layer_with_fused_normalisation_matches_normalise_then_conv():
    """Conv2d.set_input_normalization: forward and weight gradient equal those of normalising on the host first."""
    from replicann_b200.nn.layers import Conv2d
    torch.manual_seed(0)
    mean, std = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)
    a = Conv2d(3, 64, 7, 2, 3, bias=False).cuda().to(torch.bfloat16)
    b = Conv2d(3, 64, 7, 2, 3, bias=False).cuda().to(torch.bfloat16)
    b.load_state_dict(a.state_dict())
    a.set_input_normalization(mean, std)
    assert set(a.state_dict()) == set(b.state_dict())   # the constants are not part of the checkpoint
    x = rnd((4, 3, 64, 64), 9, scale=0.3) + 0.45
    xn = (x - torch.tensor(mean).view(1, 3, 1, 1)) * (1.0 / torch.tensor(std)).view(1, 3, 1, 1)
    ya, yb = a(x.cuda()), b(xn.cuda())
    assert torch.equal(ya, yb)
    dy = rnd(tuple(ya.shape), 10, torch.bfloat16).cuda().contiguous(memory_format=torch.channels_last)
    ya.backward(dy); yb.backward(dy)
    assert torch.equal(a.weight.grad, b.weight.grad)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_relu_add_cast(dtype):
    f = RF()
    x = rnd((33, 64), 1, dtype)
    y = f.relu_fwd(x.cuda())
    assert torch.equal(y.cpu(), x.relu())
    dy = rnd((33, 64), 2, dtype)
    dx = f.relu_bwd(dy.cuda(), y)
    assert torch.equal(dx.cpu(), dy * (x > 0))
    if dtype == torch.bfloat16:
        s = f.add(x.cuda(), dy.cuda())
        assert torch.equal(s.cpu(), (x.float() + dy.float()).to(torch.bfloat16))
    else:
        assert torch.equal(f.cast_f32_to_bf16(x.cuda()).cpu(), x.to(torch.bfloat16))
        assert torch.equal(f.cast_bf16_to_f32(x.to(torch.bfloat16).cuda()).cpu(), x.to(torch.bfloat16).float())
        t = f.tanh_fwd(x.cuda())
        assert rel_err(t, x.tanh()) < 1e-5
        assert rel_err(f.tanh_bwd(dy.cuda(), t), dy * (1 - x.tanh() ** 2)) < 1e-4


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_dropout_properties(dtype):
    f = RF()
    n, p = 1 << 20, 0.1
    x = torch.ones(n, dtype=dtype, device="cuda")
    y, mask = f.dropout_fwd(x, p, seed=1234, offset=0)
    keep = mask.float().mean().item()
    assert abs(keep - (1 - p)) < 5e-3                         # Bernoulli(1-p)
    kept = y[mask.bool()].float()
    assert torch.allclose(kept, torch.full_like(kept, 1 / (1 - p)), rtol=1e-2)
    assert (y[~mask.bool()] == 0).all()
    y2, mask2 = f.dropout_fwd(x, p, seed=1234, offset=0)
    assert torch.equal(mask, mask2)                           # counter-based: reproducible
    y3, mask3 = f.dropout_fwd(x, p, seed=1234, offset=1)
    assert not torch.equal(mask, mask3)
    dx = f.dropout_bwd(x, mask, p)
    assert torch.equal(dx, y)
    y0, m0 = f.dropout_fwd(x, 0.0, seed=1, offset=0)
    assert torch.equal(y0, x) and bool(m0.all())


# ------------------------------------------------------------------------------------------ fp32 path
@pytest.mark.parametrize("N,H,W,C,K,R,S,stride,pad", [(4, 28, 28, 1, 6, 5, 5, 1, 2), (4, 14, 14, 6, 16, 5, 5, 1, 0),
                                                       (2, 9, 9, 3, 8, 3, 3, 2, 1)])
def test_conv_f32(N, H, W, C, K, R, S, stride, pad):
    f = RF()
    x, w, b = rnd((N, H, W, C), 1), rnd((K, R, S, C), 2, scale=0.2), rnd((K,), 3)
    xr = x.permute(0, 3, 1, 2).contiguous().requires_grad_(True)
    wr = w.permute(0, 3, 1, 2).contiguous().requires_grad_(True)
    ref = F.conv2d(xr, wr, b, stride=stride, padding=pad)
    y = f.conv2d_fprop_f32(x.cuda(), w.cuda(), b.cuda(), stride=stride, pad=pad)
    assert rel_err(y, ref.permute(0, 2, 3, 1)) < 1e-4
    dy = rnd(tuple(y.shape), 4)
    ref.backward(dy.permute(0, 3, 1, 2))
    dx = f.conv2d_dgrad_f32(dy.cuda(), w.cuda(), tuple(x.shape), stride=stride, pad=pad)
    dw = f.conv2d_wgrad_f32(x.cuda(), dy.cuda(), tuple(w.shape), stride=stride, pad=pad)
    assert rel_err(dx, xr.grad.permute(0, 2, 3, 1)) < 1e-4
    assert rel_err(dw, wr.grad.permute(0, 2, 3, 1)) < 1e-4


@pytest.mark.parametrize("M,N,K", [(64, 120, 400), (64, 10, 84), (130, 70, 33)])
def test_gemm_f32_all_layouts(M, N, K):
    f = RF()
    a, b, bias = rnd((M, K), 1), rnd((N, K), 2), rnd((N,), 3)
    ref = a @ b.t()
    assert rel_err(f.gemm_f32(a.cuda(), b.cuda(), bias=bias.cuda()), ref + bias) < 1e-4
    assert rel_err(f.gemm_f32(a.cuda(), b.t().contiguous().cuda(), b_mn=True), ref) < 1e-4
    assert rel_err(f.gemm_f32(a.t().contiguous().cuda(), b.t().contiguous().cuda(), a_mn=True, b_mn=True), ref) < 1e-4


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-5), (torch.bfloat16, 1e-2)])
def test_colsum(dtype, tol):
    f = RF()
    x = rnd((1000, 96), 1, dtype)
    assert rel_err(f.colsum(x.cuda()), x.float().sum(0)) < tol


@pytest.mark.gpu
@pytest.mark.parametrize("M,C,dt", [(5000, 64, "bf16"), (777, 4096, "bf16"), (1030, 24, "bf16"), (999, 36, "f32"),
                                     (4096, 10, "f32")])
def test_colsum_and_fused_relu_bwd_colsum(M, C, dt):
    """Bias gradients: the vector column-sum kernel (C a multiple of the 16-byte vector) and the scalar one, and
    ReLU backward fused with the column sum, against torch on the CPU (with and without accumulation)."""
    import torch
    from replicann_b200 import functional as RF
    dtype = torch.bfloat16 if dt == "bf16" else torch.float32
    g = torch.Generator().manual_seed(M + C)
    x = torch.randn(M, C, generator=g).to(dtype)
    y = torch.randn(M, C, generator=g).to(dtype)
    ref = x.float().sum(0)
    got = RF.colsum(x.cuda())
    tol = 2e-2 if dt == "bf16" else 1e-4
    assert ((got.float().cpu() - ref).abs().max() / ref.abs().max()).item() < tol
    prev = torch.randn(C, generator=g).to(dtype)
    out = prev.clone().cuda()
    dx = RF.relu_bwd_colsum(x.cuda(), y.cuda(), out=out, accumulate=True)
    dx_ref = torch.where(y.float() > 0, x.float(), torch.zeros(()))
    assert torch.equal(dx.float().cpu(), dx_ref)                      # a select: exact
    ref2 = dx_ref.sum(0) + prev.float()
    assert ((out.float().cpu() - ref2).abs().max() / ref2.abs().max()).item() < tol


@pytest.mark.parametrize("n_out,relu", [(10, False), (100, True), (13, False)])
def test_bf16_linear_with_out_features_not_a_multiple_of_8(n_out, relu):
    """A 10- / 100-class head: the tensor-core GEMM stores 16-byte row pieces, so the layer runs zero-padded to the
    next multiple of 8; outputs and all three gradients must match torch on the unpadded shapes."""
    from replicann_b200.nn.layers import Linear
    torch.manual_seed(0)
    lin = Linear(64, n_out, relu=relu).cuda().to(torch.bfloat16)
    x = rnd((24, 64), 3, torch.bfloat16)
    dy = rnd((24, n_out), 4, torch.bfloat16)
    xr = x.float().requires_grad_(True)
    wr = lin.weight.detach().float().cpu().requires_grad_(True)
    br = lin.bias.detach().float().cpu().requires_grad_(True)
    yr = F.linear(xr, wr, br)
    yr = F.relu(yr) if relu else yr
    yr.backward(dy.float())
    xg = x.cuda().requires_grad_(True)
    y = lin(xg)
    assert tuple(y.shape) == (24, n_out)
    y.backward(dy.cuda())
    assert rel_err(y, yr) < 1e-2
    assert rel_err(xg.grad, xr.grad) < 1e-2
    assert rel_err(lin.weight.grad, wr.grad) < 1e-2 and rel_err(lin.bias.grad, br.grad) < 1e-2

