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
