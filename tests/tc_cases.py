"""Shared case definitions for the tensor-core kernel parity tests (used by pytest and tools/diag_tc.py).

Every case computes the CUDA result through the C ABI and a torch-CPU fp32 reference on the SAME
bf16-rounded inputs, and returns error statistics.  Tolerance (north_star): bf16 paths within 1e-2.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

TOL = 1e-2  # relative to max |ref| (bf16 output rounding is 2^-9 relative per element)


def _stats(got: torch.Tensor, ref: torch.Tensor) -> dict:
    got = got.detach().float().cpu()
    ref = ref.detach().float().cpu()
    err = (got - ref).abs()
    scale = ref.abs().max().item() + 1e-12
    bad = err > TOL * scale
    out = {
        "max_err": err.max().item(), "ref_max": scale, "rel": err.max().item() / scale,
        "n_bad": int(bad.sum().item()), "numel": ref.numel(), "finite": bool(torch.isfinite(got).all()),
    }
    if out["n_bad"]:
        idx = bad.nonzero()
        out["first_bad"] = idx[0].tolist()
        out["last_bad"] = idx[-1].tolist()
        # which rows / cols are affected (helps to localise descriptor mistakes)
        flat = bad.reshape(-1, bad.shape[-1])
        out["bad_rows"] = flat.any(dim=1).nonzero().flatten()[:12].tolist()
        out["bad_cols"] = flat.any(dim=0).nonzero().flatten()[:12].tolist()
    return out


def _rand(shape, seed, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(shape, generator=g) * scale).to(torch.bfloat16)


# ----------------------------------------------------------------------------------------- GEMM
GEMM_CASES = {
    # name: (M, N, K, a_mn, b_mn, bias, relu, out_f32)
    "gemm_kk_256": (256, 256, 512, False, False, False, False, False),
    "gemm_kk_ragged": (200, 136, 72, False, False, False, False, False),
    "gemm_kk_bias_relu": (384, 512, 256, False, False, True, True, False),
    "gemm_kk_bn64": (128, 64, 128, False, False, False, False, False),
    "gemm_kk_bn128": (256, 384, 192, False, False, False, False, False),
    "gemm_kk_f32": (256, 128, 256, False, False, False, False, True),
    "gemm_kk_splitk": (128, 256, 8192, False, False, True, False, False),
    "gemm_kk_big": (2048, 2048, 2048, False, False, False, False, False),
    "gemm_kk_multitile": (8192, 1024, 256, False, False, False, False, False),
    "gemm_kmn_256": (256, 256, 512, False, True, False, False, False),
    "gemm_kmn_ragged": (200, 136, 72, False, True, False, False, False),
    "gemm_mnmn_256": (256, 256, 512, True, True, False, False, False),
    "gemm_mnmn_ragged": (200, 136, 72, True, True, False, False, False),
    "gemm_mnmn_f32_splitk": (128, 128, 4096, True, True, False, False, True),
}


def run_gemm(name: str) -> dict:
    from replicann_b200 import functional as RF
    M, N, K, a_mn, b_mn, bias, relu, out_f32 = GEMM_CASES[name]
    a = _rand((M, K), 1)
    b = _rand((N, K), 2)
    bv = torch.randn(N, generator=torch.Generator().manual_seed(3)) if bias else None
    ref = a.float() @ b.float().t()
    if bv is not None:
        ref = ref + bv
    if relu:
        ref = ref.relu()
    a_dev = (a.t().contiguous() if a_mn else a).cuda()
    b_dev = (b.t().contiguous() if b_mn else b).cuda()
    d = RF.gemm(a_dev, b_dev, a_mn=a_mn, b_mn=b_mn, bias=bv.cuda() if bv is not None else None, relu=relu,
                out_dtype=torch.float32 if out_f32 else torch.bfloat16)
    torch.cuda.synchronize()
    return _stats(d, ref)


# ----------------------------------------------------------------------------------------- conv
CONV_CASES = {
    # name: (N, H, W, C, K, R, S, stride, pad)
    "conv_1x1_c64": (2, 8, 8, 64, 64, 1, 1, 1, 0),
    "conv_1x1_c128_k256": (2, 14, 14, 128, 256, 1, 1, 1, 0),
    "conv_3x3_p1": (3, 14, 14, 64, 128, 3, 3, 1, 1),
    "conv_3x3_p1_w56": (2, 56, 56, 64, 64, 3, 3, 1, 1),
    "conv_3x3_s2": (2, 28, 28, 128, 128, 3, 3, 2, 1),
    "conv_1x1_s2": (2, 28, 28, 256, 512, 1, 1, 2, 0),
    "conv_7x7_s2": (2, 32, 32, 64, 64, 7, 7, 2, 3),
    "conv_3x3_odd": (2, 7, 7, 128, 64, 3, 3, 2, 1),
    "conv_5x5_p2_k192": (2, 27, 27, 64, 192, 5, 5, 1, 2),
    # more tiles than SMs: every CTA walks several tiles (accumulator double-buffering, ring wrap-around)
    "conv_3x3_multitile": (16, 56, 56, 64, 64, 3, 3, 1, 1),
    "conv_1x1_multitile_k256": (16, 56, 56, 64, 256, 1, 1, 1, 0),
    "conv_3x3_s2_multitile": (32, 56, 56, 128, 128, 3, 3, 2, 1),
}


def _conv_inputs(name):
    N, H, W, C, K, R, S, stride, pad = CONV_CASES[name]
    x = _rand((N, H, W, C), 11)
    w = _rand((K, R, S, C), 12, scale=(R * S * C) ** -0.5)
    return x, w, stride, pad


def _ref_conv(x, w, stride, pad):
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(True)  # NCHW view
    wr = w.float().permute(0, 3, 1, 2).requires_grad_(True)  # KCRS view
    y = F.conv2d(xr, wr, stride=stride, padding=pad)
    return xr, wr, y


def run_conv_fprop(name: str, epilogue: bool = False) -> dict:
    from replicann_b200 import functional as RF
    x, w, stride, pad = _conv_inputs(name)
    _, _, y = _ref_conv(x, w, stride, pad)
    ref = y.permute(0, 2, 3, 1)
    kw = {}
    if epilogue:
        K = w.shape[0]
        g = torch.Generator().manual_seed(13)
        scale = torch.rand(K, generator=g) + 0.5
        shift = torch.randn(K, generator=g)
        res = _rand(tuple(ref.shape), 14)
        ref = (ref * scale + shift + res.float()).relu()
        kw = dict(scale=scale.cuda(), shift=shift.cuda(), residual=res.cuda(), relu=True)
    got = RF.conv2d_fprop(x.cuda(), w.cuda(), stride=stride, pad=pad, **kw)
    torch.cuda.synchronize()
    return _stats(got, ref)


def run_conv_dgrad(name: str, accumulate: bool = False) -> dict:
    from replicann_b200 import functional as RF
    x, w, stride, pad = _conv_inputs(name)
    xr, wr, y = _ref_conv(x, w, stride, pad)
    dy = _rand(tuple(y.permute(0, 2, 3, 1).shape), 15)
    y.backward(dy.float().permute(0, 3, 1, 2))
    ref = xr.grad.permute(0, 2, 3, 1)
    out = None
    if accumulate:
        base = _rand(tuple(x.shape), 16)
        ref = ref + base.float()
        out = base.cuda()
    got = RF.conv2d_dgrad(dy.cuda(), w.cuda(), tuple(x.shape), stride=stride, pad=pad, out=out,
                          accumulate=accumulate)
    torch.cuda.synchronize()
    return _stats(got, ref)


def run_conv_wgrad(name: str, f32: bool = False) -> dict:
    from replicann_b200 import functional as RF
    x, w, stride, pad = _conv_inputs(name)
    xr, wr, y = _ref_conv(x, w, stride, pad)
    dy = _rand(tuple(y.permute(0, 2, 3, 1).shape), 15)
    y.backward(dy.float().permute(0, 3, 1, 2))
    ref = wr.grad.permute(0, 2, 3, 1)
    got = RF.conv2d_wgrad(x.cuda(), dy.cuda(), tuple(w.shape), stride=stride, pad=pad,
                          out_dtype=torch.float32 if f32 else torch.bfloat16)
    torch.cuda.synchronize()
    return _stats(got, ref)


# ----------------------------------------------------------------------------------------- first layer
FIRST_CASES = {
    # name: (N, H, W, C, K, R, S, stride, pad)
    "first_7x7_s2_p3": (4, 64, 64, 3, 64, 7, 7, 2, 3),       # ResNet stem
    "first_11x11_s4_p2": (2, 67, 67, 3, 64, 11, 11, 4, 2),   # AlexNet stem (two 8-tap groups)
    "first_3x3_p1": (2, 32, 40, 3, 64, 3, 3, 1, 1),          # VGG stem
    "first_5x5_c1": (3, 28, 28, 1, 16, 5, 5, 1, 2),
    "first_7x7_multitile": (32, 224, 224, 3, 64, 7, 7, 2, 3),
}


def run_first_conv(name: str) -> dict:
    import torch.nn as nn
    from replicann_b200.nn.layers import Conv2d
    N, H, W, C, K, R, S, stride, pad = FIRST_CASES[name]
    g = torch.Generator().manual_seed(21)
    x = torch.randn((N, C, H, W), generator=g)
    ref = nn.Conv2d(C, K, (R, S), stride, pad)
    ours = Conv2d(C, K, (R, S), stride, pad, relu=True).cuda().to(torch.bfloat16)
    ours.load_state_dict(ref.state_dict())
    ref.load_state_dict({k: v.float().cpu() for k, v in ours.state_dict().items()})
    xr = x.to(torch.bfloat16).float()
    yr = F.relu(ref(xr))
    dy = _rand(tuple(yr.shape), 22).float()
    yr.backward(dy)
    yo = ours(x.cuda())
    yo.backward(dy.cuda().to(torch.bfloat16).contiguous(memory_format=torch.channels_last))
    torch.cuda.synchronize()
    st = _stats(yo, yr)
    sw = _stats(ours.weight.grad, ref.weight.grad)
    sb = _stats(ours.bias.grad, ref.bias.grad)
    st["wgrad_rel"], st["wgrad_bad"] = sw["rel"], sw["n_bad"]
    st["bgrad_rel"] = sb["rel"]
    st["n_bad"] += sw["n_bad"] + sb["n_bad"]
    return st
