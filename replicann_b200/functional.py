"""Thin tensor-level wrappers over the C ABI (one Python function per exported kernel entry point).

Tensors are torch CUDA tensors used purely as device buffers; layouts are the library's own:
activations NHWC, filters KRSC, everything contiguous.  No function here computes on the host.
"""
from __future__ import annotations

import torch

from . import _lib
from ._lib import check, conv_desc, cur_stream, ptr

_workspaces: dict[int, torch.Tensor] = {}


def workspace(nbytes: int, device) -> torch.Tensor | None:
    """Grow-only per-device scratch buffer (torch caching allocator owns the memory)."""
    if nbytes <= 0:
        return None
    idx = torch.device(device).index or 0
    ws = _workspaces.get(idx)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=device)
        _workspaces[idx] = ws
    return ws


def _req(t: torch.Tensor, dtype, name: str):
    if not t.is_cuda:
        raise _lib.RnnError(f"{name} must be a CUDA tensor: replicann_b200 has no CPU path")
    if t.dtype != dtype:
        raise TypeError(f"{name} must be {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise ValueError(f"{name} must be contiguous")


# ------------------------------------------------------------------------------------------- GEMM
def gemm(a: torch.Tensor, b: torch.Tensor, *, a_mn: bool = False, b_mn: bool = False,
         bias: torch.Tensor | None = None, relu: bool = False, out_dtype=torch.bfloat16) -> torch.Tensor:
    """D[M,N] = op(A) . op(B)^T.  a_mn: A stored [K,M]; b_mn: B stored [K,N]."""
    lib = _lib.load()
    _req(a, torch.bfloat16, "a")
    _req(b, torch.bfloat16, "b")
    K, M = (a.shape if a_mn else a.shape[::-1])
    Kb, N = (b.shape if b_mn else b.shape[::-1])
    if K != Kb:
        raise ValueError(f"reduction mismatch {K} vs {Kb}")
    out_f32 = out_dtype == torch.float32
    d = torch.empty((M, N), dtype=out_dtype, device=a.device)
    if bias is not None:
        _req(bias, torch.float32, "bias")
    nbytes = lib.rnn_gemm_bf16_workspace(M, N, K, int(out_f32), int(bias is not None or relu))
    ws = workspace(nbytes, a.device)
    check(lib.rnn_gemm_bf16(ptr(a), ptr(b), ptr(d), M, N, K, int(a_mn), int(b_mn), ptr(bias), int(relu),
                            int(out_f32), ptr(ws), nbytes, cur_stream()), "rnn_gemm_bf16")
    return d


# ------------------------------------------------------------------------------------------- conv
def _cdesc(x_shape, w_shape, stride, pad):
    N, H, W, Cin = x_shape
    K, R, S, C2 = w_shape
    if Cin != C2:
        raise ValueError(f"channel mismatch {Cin} vs {C2}")
    return conv_desc(N, H, W, Cin, K, R, S, stride, pad)


def conv2d_fprop(x, w, *, stride=1, pad=0, scale=None, shift=None, residual=None, relu=False):
    """x [N,H,W,C] bf16, w [K,R,S,C] bf16 -> y [N,P,Q,K] bf16."""
    lib = _lib.load()
    _req(x, torch.bfloat16, "x")
    _req(w, torch.bfloat16, "w")
    d = _cdesc(x.shape, w.shape, stride, pad)
    y = torch.empty((d.N, d.P, d.Q, d.K), dtype=torch.bfloat16, device=x.device)
    if residual is not None:
        _req(residual, torch.bfloat16, "residual")
    check(lib.rnn_conv2d_fprop_bf16(d, ptr(x), ptr(w), ptr(y), ptr(scale), ptr(shift), ptr(residual), int(relu),
                                    cur_stream()), "rnn_conv2d_fprop_bf16")
    return y


def conv2d_dgrad(dy, w, x_shape, *, stride=1, pad=0, out=None, accumulate=False):
    lib = _lib.load()
    _req(dy, torch.bfloat16, "dy")
    _req(w, torch.bfloat16, "w")
    d = _cdesc(x_shape, w.shape, stride, pad)
    if tuple(dy.shape) != (d.N, d.P, d.Q, d.K):
        raise ValueError(f"dy shape {tuple(dy.shape)} != {(d.N, d.P, d.Q, d.K)}")
    dx = out if out is not None else torch.empty(tuple(x_shape), dtype=torch.bfloat16, device=dy.device)
    check(lib.rnn_conv2d_dgrad_bf16(d, ptr(dy), ptr(w), ptr(dx), int(accumulate), cur_stream()),
          "rnn_conv2d_dgrad_bf16")
    return dx


def conv2d_wgrad(x, dy, w_shape, *, stride=1, pad=0, out=None, accumulate=False, out_dtype=torch.bfloat16):
    lib = _lib.load()
    _req(x, torch.bfloat16, "x")
    _req(dy, torch.bfloat16, "dy")
    d = _cdesc(x.shape, w_shape, stride, pad)
    dw = out if out is not None else torch.empty(tuple(w_shape), dtype=out_dtype, device=x.device)
    nbytes = lib.rnn_conv2d_wgrad_workspace(d)
    ws = workspace(nbytes, x.device)
    check(lib.rnn_conv2d_wgrad_bf16(d, ptr(x), ptr(dy), ptr(dw), int(dw.dtype == torch.float32), int(accumulate),
                                    ptr(ws), nbytes, cur_stream()), "rnn_conv2d_wgrad_bf16")
    return dw
