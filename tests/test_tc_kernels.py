"""GPU parity of the tcgen05 GEMM / implicit-GEMM convolution kernels (through the C ABI) against
torch CPU fp32 on identical bf16-rounded inputs."""
import pytest

import tc_cases as T

pytestmark = pytest.mark.gpu


def _ok(st):
    assert st["finite"], st
    assert st["n_bad"] == 0, st


@pytest.mark.parametrize("name", list(T.GEMM_CASES))
def test_gemm(name):
    _ok(T.run_gemm(name))


@pytest.mark.parametrize("name", list(T.CONV_CASES))
def test_conv_fprop(name):
    _ok(T.run_conv_fprop(name))


@pytest.mark.parametrize("name", ["conv_3x3_p1", "conv_1x1_c128_k256"])
def test_conv_fprop_epilogue(name):
    _ok(T.run_conv_fprop(name, epilogue=True))


@pytest.mark.parametrize("name", list(T.CONV_CASES))
def test_conv_dgrad(name):
    _ok(T.run_conv_dgrad(name))


def test_conv_dgrad_accumulate():
    _ok(T.run_conv_dgrad("conv_1x1_s2", accumulate=True))
    _ok(T.run_conv_dgrad("conv_3x3_p1", accumulate=True))


@pytest.mark.parametrize("name", list(T.CONV_CASES))
def test_conv_wgrad(name):
    _ok(T.run_conv_wgrad(name))


def test_conv_wgrad_f32():
    _ok(T.run_conv_wgrad("conv_3x3_p1", f32=True))


@pytest.mark.parametrize("name", list(T.FIRST_CASES))
def test_first_layer_packed_pixel_conv(name):
    _ok(T.run_first_conv(name))


@pytest.mark.parametrize("name", ["conv_1x1_c64", "conv_1x1_c128_k256", "conv_3x3_p1", "conv_1x1_s2", "conv_3x3_multitile",
                                  "conv_1x1_multitile_k256"])
def test_conv_fprop_with_fused_bn_statistics(name):
    """The epilogue statistics must equal the separate statistics kernel run on the stored output."""
    import torch
    from replicann_b200 import functional as RF
    x, w, stride, pad = T._conv_inputs(name)
    K = w.shape[0]
    g = torch.Generator().manual_seed(3)
    gamma, beta = (torch.rand(K, generator=g) + 0.5).cuda(), torch.randn(K, generator=g).cuda()
    rm1, rv1 = torch.zeros(K, device="cuda"), torch.ones(K, device="cuda")
    rm2, rv2 = torch.zeros(K, device="cuda"), torch.ones(K, device="cuda")
    y1, *st1 = RF.conv2d_fprop_bn_stats(x.cuda(), w.cuda(), gamma, beta, 1e-5, 0.1, rm1, rv1, stride=stride, pad=pad)
    y2 = RF.conv2d_fprop(x.cuda(), w.cuda(), stride=stride, pad=pad)
    st2 = RF.bn_stats(y2.reshape(-1, K), gamma, beta, 1e-5, 0.1, rm2, rv2)
    assert torch.equal(y1, y2)
    for a, b, nm in zip(st1, st2, ("mean", "invstd", "scale", "shift")):
        rel = ((a - b).abs().max() / (b.abs().max() + 1e-12)).item()
        assert rel < 2e-5, (nm, rel)
    assert torch.allclose(rm1, rm2, rtol=1e-5, atol=1e-7) and torch.allclose(rv1, rv2, rtol=1e-5, atol=1e-7)


@pytest.mark.gpu
@pytest.mark.parametrize("M,N,K,a_mn,b_mn", [(512, 256, 384, False, False), (1000, 264, 516, False, True),
                                              (256, 512, 1024, True, True), (8192, 512, 512, False, False)])
def test_gemm_f32_on_tensor_cores_is_fp32_accurate(M, N, K, a_mn, b_mn):
    """rnn_gemm_f32_tc: fp32 operands split into three bf16 pieces, six piece products as one tcgen05 GEMM.
    Against a float64 CPU product the error must be at fp32 level (1e-5 of max|ref|; a plain bf16 GEMM is 1e-2,
    TF32 1e-3), with bias + ReLU fused."""
    import torch
    from replicann_b200 import functional as RF
    g = torch.Generator().manual_seed(M + N + K)
    a = torch.randn((K, M) if a_mn else (M, K), generator=g)
    b = torch.randn((K, N) if b_mn else (N, K), generator=g)
    bias = torch.randn(N, generator=g)
    ref = (a.double().t() if a_mn else a.double()) @ (b.double() if b_mn else b.double().t())
    got = RF.gemm_f32(a.cuda(), b.cuda(), a_mn=a_mn, b_mn=b_mn)
    err = (got.cpu().double() - ref).abs().max().item() / ref.abs().max().item()
    assert err < 1e-5, err
    ref2 = torch.relu(ref + bias.double())
    got2 = RF.gemm_f32(a.cuda(), b.cuda(), a_mn=a_mn, b_mn=b_mn, bias=bias.cuda(), relu=True)
    err2 = (got2.cpu().double() - ref2).abs().max().item() / ref2.abs().max().item()
    assert err2 < 1e-5, err2


@pytest.mark.parametrize("N,H,W,C,K,R,pad", [(4, 28, 28, 64, 64, 3, 1), (2, 30, 17, 128, 192, 3, 1), (3, 14, 14, 256, 256, 1, 0),
                                             (2, 56, 56, 64, 128, 3, 1)])
def test_relu_mask_from_the_fprop_epilogue_and_into_the_dgrad_epilogue(N, H, W, C, K, R, pad):
    """rnn_conv2d_fprop_relu_mask_bf16 / rnn_conv2d_dgrad_masked_bf16: y is bit-identical to the ReLU-fused fprop, the
    mask is exactly (y > 0) packed 8 channels per byte, and the masked dgrad equals dgrad followed by ReLU backward
    (bit for bit) -- and the torch gradient of conv(relu(.)) within the bf16 tolerance."""
    import torch
    import torch.nn.functional as F
    from replicann_b200 import functional as RF
    g = torch.Generator().manual_seed(3)
    x = torch.randn(N, H, W, C, generator=g).to(torch.bfloat16).cuda()
    w = (torch.randn(K, R, R, C, generator=g) * 0.05).to(torch.bfloat16).cuda()
    b = torch.randn(K, generator=g).cuda()
    y_ref = RF.conv2d_fprop(x, w, stride=1, pad=pad, shift=b, relu=True)
    y, mask = RF.conv2d_fprop_relu_mask(x, w, stride=1, pad=pad, shift=b)
    assert torch.equal(y, y_ref)
    bits = (y > 0).reshape(N, y.shape[1], y.shape[2], K // 8, 8).to(torch.int32)
    packed = (bits * (1 << torch.arange(8, device="cuda", dtype=torch.int32))).sum(-1).to(torch.uint8)
    assert torch.equal(mask, packed)
    # consumer: a second convolution over y; its dgrad with the mask == dgrad then relu_bwd
    w2 = (torch.randn(64, 3, 3, K, generator=g) * 0.05).to(torch.bfloat16).cuda()
    dy2 = torch.randn(N, y.shape[1], y.shape[2], 64, generator=g).to(torch.bfloat16).cuda()
    dx_plain = RF.conv2d_dgrad(dy2, w2, tuple(y.shape), stride=1, pad=1)
    dx_ref = RF.relu_bwd(dx_plain, y)
    dx = RF.conv2d_dgrad(dy2, w2, tuple(y.shape), stride=1, pad=1, relu_mask=mask)
    assert torch.equal(dx, dx_ref)
    pre = torch.where(y > 0, y.float(), torch.full_like(y.float(), -1.0)).cpu().permute(0, 3, 1, 2).requires_grad_(True)
    out = F.conv2d(torch.relu(pre), w2.float().cpu().permute(0, 3, 1, 2), padding=1)
    out.backward(dy2.float().cpu().permute(0, 3, 1, 2))
    ref = pre.grad.permute(0, 2, 3, 1)
    err = (dx.float().cpu() - ref).abs().max() / (ref.abs().max() + 1e-20)
    assert err < 1e-2, err
