"""Tensor-core kernels at the shapes bench.py actually launches (BASELINE.json configs[1..4] at full batch).

At N = 256 the library picks launch forms the small parity cases never reach: CTA pairs, split factors of the
weight gradient, persistent CTAs that wrap the two-deep TMEM ring many times, the forced grid of the
statistics epilogue.  Each case runs fprop (+ fused BatchNorm statistics), dgrad and wgrad through the C ABI on
the full tensors and compares them with torch CPU fp32 on the same bf16-rounded inputs.  The CPU reference is
evaluated in chunks of images (convolutions are independent per image; the filter gradient is a sum over
chunks), so the host never holds more than one chunk of fp32 activations.

Tolerance (BASELINE.json north_star): bf16 paths within 1e-2 of max|ref| per element.
"""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

TOL = 1e-2

REAL_CONV = {
    # name: (N, H, W, C, K, R, S, stride, pad, chunk)
    "r50_s1_3x3_c64": (256, 56, 56, 64, 64, 3, 3, 1, 1, 32),          # C64 3x3 56^2 (worst roofline shape)
    "r50_s1_1x1_c64_k256": (256, 56, 56, 64, 256, 1, 1, 1, 0, 32),    # + statistics epilogue
    "r50_s2_3x3_c128_s2": (256, 56, 56, 128, 128, 3, 3, 2, 1, 32),    # strided dgrad (4 parity classes)
    "r50_s4_1x1_c2048_k512": (256, 7, 7, 2048, 512, 1, 1, 1, 0, 64),  # late stage, 1.3 waves
    "vgg_3x3_c64_224": (128, 224, 224, 64, 64, 3, 3, 1, 1, 4),        # VGG-16 conv1_2 at batch 128
}


def _rand(shape, seed, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(shape, generator=g) * scale).to(torch.bfloat16)


class _Err:
    """Running max-error statistics of a tensor compared chunk by chunk."""

    def __init__(self):
        self.err, self.ref = 0.0, 0.0
        self.finite = True

    def add(self, got, ref):
        got = got.detach().float().cpu()
        self.finite &= bool(torch.isfinite(got).all())
        self.err = max(self.err, (got - ref).abs().max().item())
        self.ref = max(self.ref, ref.abs().max().item())

    @property
    def rel(self):
        return self.err / (self.ref + 1e-12)


@pytest.mark.parametrize("name", list(REAL_CONV))
def test_conv_passes_at_bench_shape(name):
    from replicann_b200 import functional as RF
    N, H, W, C, K, R, S, stride, pad, chunk = REAL_CONV[name]
    x = _rand((N, H, W, C), 11)
    w = _rand((K, R, S, C), 12, scale=(R * S * C) ** -0.5)
    P = (H + 2 * pad - R) // stride + 1
    Q = (W + 2 * pad - S) // stride + 1
    dy = _rand((N, P, Q, K), 15)
    xg, wg, dyg = x.cuda(), w.cuda(), dy.cuda()

    # GPU: all three passes on the full tensors, plus the statistics-epilogue form of fprop
    y = RF.conv2d_fprop(xg, wg, stride=stride, pad=pad)
    gamma = torch.ones(K, device="cuda")
    beta = torch.zeros(K, device="cuda")
    rm, rv = torch.zeros(K, device="cuda"), torch.ones(K, device="cuda")
    y_st, mean, invstd, _, _ = RF.conv2d_fprop_bn_stats(xg, wg, gamma, beta, 1e-5, 0.1, rm, rv, stride=stride, pad=pad)
    dx = RF.conv2d_dgrad(dyg, wg, (N, H, W, C), stride=stride, pad=pad)
    dw = RF.conv2d_wgrad(xg, dyg, (K, R, S, C), stride=stride, pad=pad, out_dtype=torch.float32)
    torch.cuda.synchronize()
    assert torch.equal(y, y_st)   # the statistics epilogue must not change what is stored

    # CPU reference, chunked over images
    wr = w.float().permute(0, 3, 1, 2).contiguous()
    e_y, e_dx = _Err(), _Err()
    dw_ref = torch.zeros((K, C, R, S), dtype=torch.float64)
    s1 = torch.zeros(K, dtype=torch.float64)
    s2 = torch.zeros(K, dtype=torch.float64)
    for i in range(0, N, chunk):
        xr = x[i:i + chunk].float().permute(0, 3, 1, 2).requires_grad_(True)
        wl = wr.clone().requires_grad_(True)
        yr = F.conv2d(xr, wl, stride=stride, padding=pad)
        yr.backward(dy[i:i + chunk].float().permute(0, 3, 1, 2))
        e_y.add(y[i:i + chunk], yr.detach().permute(0, 2, 3, 1))
        e_dx.add(dx[i:i + chunk], xr.grad.permute(0, 2, 3, 1))
        dw_ref += wl.grad.double()
        yq = y[i:i + chunk].float().cpu().double().reshape(-1, K)   # statistics are those of the STORED bf16 values
        s1 += yq.sum(0)
        s2 += (yq * yq).sum(0)
    e_dw = _Err()
    e_dw.add(dw, dw_ref.float().permute(0, 2, 3, 1))
    cnt = N * P * Q
    mean_ref = s1 / cnt
    var_ref = (s2 / cnt - mean_ref * mean_ref).clamp_min(0)
    invstd_ref = (var_ref + 1e-5).rsqrt()
    mean_err = ((mean.double().cpu() - mean_ref).abs().max() / (var_ref.sqrt().max() + 1e-12)).item()
    is_err = ((invstd.double().cpu() - invstd_ref).abs() / invstd_ref).max().item()
    print(f"STAT {name}: fprop {e_y.rel:.2e} dgrad {e_dx.rel:.2e} wgrad {e_dw.rel:.2e} "
          f"mean {mean_err:.2e} invstd {is_err:.2e}", flush=True)
    for e, what in ((e_y, "fprop"), (e_dx, "dgrad"), (e_dw, "wgrad")):
        assert e.finite, what
        assert e.rel < TOL, (what, e.rel)
    assert mean_err < 1e-3 and is_err < 1e-3, (mean_err, is_err)


def test_alexnet_fc_gemm_at_batch_256():
    """AlexNet fc1 at batch 256 (configs[1]): M256 x K9216 x N4096, forward (bias + ReLU), dgrad and wgrad forms."""
    from replicann_b200 import functional as RF
    M, K, N = 256, 9216, 4096
    a = _rand((M, K), 1)
    w = _rand((N, K), 2, scale=K ** -0.5)
    bias = torch.randn(N, generator=torch.Generator().manual_seed(3))
    dy = _rand((M, N), 4)
    ag, wg, dyg = a.cuda(), w.cuda(), dy.cuda()
    y = RF.gemm(ag, wg, bias=bias.cuda(), relu=True)                      # y = relu(a w^T + b)
    dx = RF.gemm(dyg, wg, b_mn=True)                                      # dx = dy w
    dw = RF.gemm(dyg, ag, a_mn=True, b_mn=True, out_dtype=torch.float32)  # dw = dy^T a
    torch.cuda.synchronize()
    af, wf, dyf = a.float(), w.float(), dy.float()
    for got, ref, what in ((y, (af @ wf.t() + bias).relu(), "fwd"), (dx, dyf @ wf, "dgrad"), (dw, dyf.t() @ af, "wgrad")):
        e = _Err()
        e.add(got, ref)
        print(f"STAT alexnet_fc1 {what}: {e.rel:.2e}", flush=True)
        assert e.finite and e.rel < TOL, (what, e.rel)
