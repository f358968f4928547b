"""Thin tensor-level wrappers over the C ABI (one Python function per exported kernel entry point).

Tensors are torch CUDA tensors used purely as device buffers; layouts are the library's own:
activations NHWC, filters KRSC, everything contiguous.  No function here computes on the host.
"""
from __future__ import annotations

import contextlib
import os
import threading

import torch

from . import _lib
from ._lib import check, conv_desc, cur_stream, ptr

_workspaces: dict[tuple[int, int], torch.Tensor] = {}
_gemm_ws_bytes: dict[tuple, int] = {}   # workspace size per GEMM shape (saves one library call per launch)
_tls = threading.local()

# Optional per-call device timing used by bench.py: when `_profile` is a list, every instrumented wrapper
# appends (kind, algorithmic_bytes_or_None, start_event, end_event) recorded on the launching stream.
_profile: list | None = None


def set_profile(sink: list | None) -> None:
    global _profile
    _profile = sink


class _span:
    __slots__ = ("kind", "work", "e0", "tag")

    def __init__(self, kind: str, work=None, tag: str = ""):
        self.kind, self.work, self.e0, self.tag = kind, work, None, tag

    def __enter__(self):
        if _profile is not None:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e0.record()
        return self

    def __exit__(self, *exc):
        if self.e0 is not None and _profile is not None:
            e1 = torch.cuda.Event(enable_timing=True)
            e1.record()
            _profile.append((self.kind, self.work, self.e0, e1, self.tag))
        return False


def timing_enable(on: bool) -> None:
    """Library-side CUDA-event pairs around individual tensor-core / split-K / finalize launches (bench.py)."""
    check(_lib.load().rnn_timing_enable(int(on)), "rnn_timing_enable")


def timing_read(family: str) -> tuple[float, int]:
    """(summed device ms, launches) of one family since timing_enable(True); waits for the recorded events."""
    import ctypes as C
    ms, n = C.c_float(), C.c_int()
    check(_lib.load().rnn_timing_read(family.encode(), C.byref(ms), C.byref(n)), "rnn_timing_read")
    return float(ms.value), int(n.value)


def rng_advance() -> None:
    """Advance the device-side dropout epoch on the current stream (see rnn_rng_advance in the header): call it at
    the head of a step that is being captured into a CUDA graph, so that every replay draws new dropout masks."""
    check(_lib.load().rnn_rng_advance(cur_stream()), "rnn_rng_advance")


def launch_count() -> int:
    return int(_lib.load().rnn_launch_count())


def workspace(nbytes: int, device) -> torch.Tensor | None:
    """Grow-only per-device scratch buffer (torch caching allocator owns the memory).  Kernels enqueued inside
    `side_stream()` get a buffer of their own: they run concurrently with the main stream's kernels."""
    if nbytes <= 0:
        return None
    key = (torch.device(device).index or 0, getattr(_tls, "slot", 0))
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=device)
        _workspaces[key] = ws
    return ws


# ------------------------------------------------------------------------------------------- side stream
# Backward has two independent consumers of every convolution's output gradient: dgrad (on the critical path to
# the next layer) and wgrad (needed only by the optimizer).  The weight-gradient kernels are tensor-core work, the
# BatchNorm backward kernels that follow dgrad on the critical path are HBM streaming work; issued on two streams
# they share the SMs instead of taking turns.  RNN_SIDE_STREAM=0 keeps everything on one stream.
class _Side:
    def __init__(self, device):
        self.stream = torch.cuda.Stream(device=device)
        self.pending: list = []        # [(event after the group's last kernel, tensors kept alive, on_join)]


_sides: dict[int, _Side] = {}
_side_enabled = os.environ.get("RNN_SIDE_STREAM", "1") != "0"


def side_enabled() -> bool:
    return _side_enabled and _profile is None   # per-kernel event timing needs one stream


def _side(device) -> _Side:
    idx = torch.device(device).index
    if idx is None:
        idx = torch.cuda.current_device()
    st = _sides.get(idx)
    if st is None:
        st = _sides[idx] = _Side(torch.device("cuda", idx))
    return st


@contextlib.contextmanager
def side_stream(device):
    """Enqueue the enclosed kernels on the device's side stream, ordered after everything enqueued so far on the
    current stream.  Follow with side_commit(); the work is joined back by side_join()."""
    st = _side(device)
    ev = torch.cuda.Event()
    ev.record()
    st.stream.wait_event(ev)
    prev = getattr(_tls, "slot", 0)
    _tls.slot = 1
    try:
        with torch.cuda.stream(st.stream):
            yield st
    finally:
        _tls.slot = prev


def side_commit(device, keepalive, on_join=None) -> None:
    """Close a group of side-stream kernels.  `keepalive`: every tensor those kernels read or write that could
    otherwise be freed (and its memory reused by the main stream) before the group is joined."""
    st = _side(device)
    ev = torch.cuda.Event()
    ev.record(st.stream)
    st.pending.append((ev, keepalive, on_join))


def side_join(device=None, keep: int = 0) -> None:
    """Make the current stream wait for all but the newest `keep` committed groups, then run their on_join
    callbacks (gradient-ready notifications) and drop the tensors they kept alive."""
    for idx, st in list(_sides.items()):
        if device is not None and (torch.device(device).index or 0) != idx:
            continue
        while len(st.pending) > keep:
            ev, _keep, cb = st.pending.pop(0)
            torch.cuda.current_stream(torch.device("cuda", idx)).wait_event(ev)
            if cb is not None:
                cb()


def side_join_at_end_of_backward(device) -> None:
    """Queue an autograd-engine callback that joins every outstanding group when the running backward pass ends,
    so that whatever comes after `.backward()` -- optimizer, all-reduce, the end of a CUDA-graph capture -- sees
    the weight gradients.  Queued by every block (the callback is a no-op once nothing is pending): no state
    survives a backward pass that raised half-way."""
    dev = torch.device(device)

    def _cb():
        with torch.cuda.device(dev):
            side_join(dev)

    torch.autograd.Variable._execution_engine.queue_callback(_cb)


def _req(t: torch.Tensor, dtype, name: str):
    if not t.is_cuda:
        raise _lib.RnnError(f"{name} must be a CUDA tensor: replicann_b200 has no CPU path")
    if t.dtype != dtype:
        raise TypeError(f"{name} must be {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise ValueError(f"{name} must be contiguous")


# ------------------------------------------------------------------------------------------- GEMM
def gemm(a: torch.Tensor, b: torch.Tensor, *, a_mn: bool = False, b_mn: bool = False,
         bias: torch.Tensor | None = None, relu: bool = False, out_dtype=torch.bfloat16,
         out: torch.Tensor | None = None) -> torch.Tensor:
    """D[M,N] = op(A) . op(B)^T.  a_mn: A stored [K,M]; b_mn: B stored [K,N]."""
    lib = _lib.load()
    _req(a, torch.bfloat16, "a")
    _req(b, torch.bfloat16, "b")
    K, M = (a.shape if a_mn else a.shape[::-1])
    Kb, N = (b.shape if b_mn else b.shape[::-1])
    if K != Kb:
        raise ValueError(f"reduction mismatch {K} vs {Kb}")
    if out is not None:
        if tuple(out.shape) != (M, N) or not out.is_contiguous():
            raise ValueError(f"out must be a contiguous [{M},{N}] tensor")
        out_dtype = out.dtype
    out_f32 = out_dtype == torch.float32
    d = out if out is not None else torch.empty((M, N), dtype=out_dtype, device=a.device)
    if bias is not None:
        _req(bias, torch.float32, "bias")
    key = (M, N, K, out_f32, bias is not None or relu)
    nbytes = _gemm_ws_bytes.get(key)
    if nbytes is None:
        nbytes = _gemm_ws_bytes[key] = lib.rnn_gemm_bf16_workspace(M, N, K, int(out_f32), int(key[4]))
    ws = workspace(nbytes, a.device)
    with _span("tc_gemm", 2.0 * M * N * K, f"M{M} N{N} K{K} a_mn{int(a_mn)} b_mn{int(b_mn)}"):
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


def _conv_flops(d) -> float:
    return 2.0 * d.N * d.P * d.Q * d.K * d.C * d.R * d.S


def _conv_tag(d) -> str:
    return f"N{d.N} {d.H}x{d.W} C{d.C} K{d.K} {d.R}x{d.S} s{d.stride_h} p{d.pad_h}"


def conv2d_fprop(x, w, *, stride=1, pad=0, scale=None, shift=None, residual=None, relu=False):
    """x [N,H,W,C] bf16, w [K,R,S,C] bf16 -> y [N,P,Q,K] bf16."""
    lib = _lib.load()
    _req(x, torch.bfloat16, "x")
    _req(w, torch.bfloat16, "w")
    d = _cdesc(x.shape, w.shape, stride, pad)
    y = torch.empty((d.N, d.P, d.Q, d.K), dtype=torch.bfloat16, device=x.device)
    if residual is not None:
        _req(residual, torch.bfloat16, "residual")
    with _span("tc_conv_fprop", _conv_flops(d), _conv_tag(d)):
        check(lib.rnn_conv2d_fprop_bf16(d, ptr(x), ptr(w), ptr(y), ptr(scale), ptr(shift), ptr(residual),
                                        int(relu), cur_stream()), "rnn_conv2d_fprop_bf16")
    return y


def conv2d_fprop_relu_mask(x, w, *, stride=1, pad=0, shift=None):
    """y = relu(conv(x, w) + shift) and the 1-bit mask (y > 0), [N,P,Q,K/8] uint8, that the consumer's
    conv2d_dgrad(relu_mask=...) applies in place of a ReLU backward pass.  K % 64 == 0."""
    lib = _lib.load()
    _req(x, torch.bfloat16, "x")
    _req(w, torch.bfloat16, "w")
    d = _cdesc(x.shape, w.shape, stride, pad)
    y = torch.empty((d.N, d.P, d.Q, d.K), dtype=torch.bfloat16, device=x.device)
    mask = torch.empty((d.N, d.P, d.Q, d.K // 8), dtype=torch.uint8, device=x.device)
    with _span("tc_conv_fprop", _conv_flops(d), _conv_tag(d)):
        check(lib.rnn_conv2d_fprop_relu_mask_bf16(d, ptr(x), ptr(w), ptr(y), ptr(shift), ptr(mask), cur_stream()),
              "rnn_conv2d_fprop_relu_mask_bf16")
    return y, mask


def conv2d_fprop_bn_stats(x, w, gamma, beta, eps, momentum, running_mean, running_var, *, stride=1, pad=0):
    """Convolution whose epilogue also produces the BatchNorm batch statistics of its output.
    Returns (y, mean, invstd, scale, shift); updates the running statistics in place."""
    lib = _lib.load()
    _req(x, torch.bfloat16, "x"); _req(w, torch.bfloat16, "w")
    d = _cdesc(x.shape, w.shape, stride, pad)
    y = torch.empty((d.N, d.P, d.Q, d.K), dtype=torch.bfloat16, device=x.device)
    st = torch.empty((4, d.K), dtype=torch.float32, device=x.device)
    nbytes = lib.rnn_conv2d_fprop_bn_stats_workspace(d)
    ws = workspace(nbytes, x.device)
    with _span("tc_conv_fprop", _conv_flops(d), "stats " + _conv_tag(d)):
        check(lib.rnn_conv2d_fprop_bn_stats_bf16(d, ptr(x), ptr(w), ptr(y), ptr(gamma), ptr(beta), float(eps),
                                                 float(momentum), ptr(running_mean), ptr(running_var), ptr(st[0]),
                                                 ptr(st[1]), ptr(st[2]), ptr(st[3]), ptr(ws), nbytes, cur_stream()),
              "rnn_conv2d_fprop_bn_stats_bf16")
    return y, st[0], st[1], st[2], st[3]


# ---- first layer (C <= 8): packed-pixel layout -------------------------------------------------------
def first_conv_geometry(d):
    import ctypes as C
    hp, wp, sg = C.c_int(), C.c_int(), C.c_int()
    check(_lib.load().rnn_conv2d_first_geometry(d, C.byref(hp), C.byref(wp), C.byref(sg)), "rnn_conv2d_first_geometry")
    return hp.value, wp.value, sg.value


def _norm_args(mean, inv_std, Cn):
    if (mean is None) != (inv_std is None):
        raise ValueError("give both mean and inv_std, or neither")
    if mean is not None:
        _req(mean, torch.float32, "mean"); _req(inv_std, torch.float32, "inv_std")
        if mean.numel() != Cn or inv_std.numel() != Cn:
            raise ValueError(f"mean / inv_std must have {Cn} entries (one per input channel)")


def pack_first_input(x_nchw_f32, d, mean=None, inv_std=None):
    """fp32 NCHW images -> bf16 [N, Hp, Wp, 8], spatially zero-padded, channels padded to 8; optional per-channel
    normalisation (x - mean) * inv_std in the same pass."""
    lib = _lib.load()
    _req(x_nchw_f32, torch.float32, "x")
    hp, wp, _ = first_conv_geometry(d)
    N, Cn, H, W = x_nchw_f32.shape
    _norm_args(mean, inv_std, Cn)
    y = torch.empty((N, hp, wp, 8), dtype=torch.bfloat16, device=x_nchw_f32.device)
    with _span("layout", 4.0 * N * Cn * H * W + 16.0 * N * hp * wp, f"N{N} C{Cn} {H}x{W} -> {hp}x{wp}x8"):
        check(lib.rnn_nchw_f32_to_nhwc_bf16_padded_norm(ptr(x_nchw_f32), ptr(y), N, Cn, H, W, 8, d.pad_h, d.pad_w, hp,
                                                        wp, ptr(mean), ptr(inv_std), cur_stream()),
              "rnn_nchw_f32_to_nhwc_bf16_padded_norm")
    return y


def first_conv_fprop(x_packed, w_packed, d, *, scale=None, shift=None, relu=False, relu_mask=False):
    """relu_mask=True: y = relu(conv + shift) and the 1-bit mask (see conv2d_fprop_relu_mask); returns (y, mask)."""
    lib = _lib.load()
    _req(x_packed, torch.bfloat16, "x_packed"); _req(w_packed, torch.bfloat16, "w_packed")
    y = torch.empty((d.N, d.P, d.Q, d.K), dtype=torch.bfloat16, device=x_packed.device)
    if relu_mask:
        mask = torch.empty((d.N, d.P, d.Q, d.K // 8), dtype=torch.uint8, device=x_packed.device)
        with _span("tc_conv_fprop", _conv_flops(d), "first " + _conv_tag(d)):
            check(lib.rnn_conv2d_first_fprop_relu_mask_bf16(d, ptr(x_packed), ptr(w_packed), ptr(y), ptr(shift),
                                                            ptr(mask), cur_stream()),
                  "rnn_conv2d_first_fprop_relu_mask_bf16")
        return y, mask
    with _span("tc_conv_fprop", _conv_flops(d), "first " + _conv_tag(d)):
        check(lib.rnn_conv2d_first_fprop_bf16(d, ptr(x_packed), ptr(w_packed), ptr(y), ptr(scale), ptr(shift), int(relu),
                                              cur_stream()), "rnn_conv2d_first_fprop_bf16")
    return y


def first_conv_wgrad(x_packed, dy, d, sg, *, out_dtype=torch.bfloat16):
    lib = _lib.load()
    _req(x_packed, torch.bfloat16, "x_packed"); _req(dy, torch.bfloat16, "dy")
    dw = torch.empty((d.K, d.R, sg, 64), dtype=out_dtype, device=dy.device)
    nbytes = lib.rnn_conv2d_first_wgrad_workspace(d)
    ws = workspace(nbytes, dy.device)
    with _span("tc_conv_wgrad", _conv_flops(d), "first " + _conv_tag(d)):
        check(lib.rnn_conv2d_first_wgrad_bf16(d, ptr(x_packed), ptr(dy), ptr(dw), int(out_dtype == torch.float32),
                                              ptr(ws), nbytes, cur_stream()), "rnn_conv2d_first_wgrad_bf16")
    return dw


def conv2d_dgrad(dy, w, x_shape, *, stride=1, pad=0, out=None, accumulate=False, relu_mask=None):
    """relu_mask: the 1-bit mask of the ReLU that produced this convolution's input (conv2d_fprop_relu_mask of the
    layer before); its backward is applied in the epilogue (unit stride, no accumulation, C % 64 == 0)."""
    lib = _lib.load()
    _req(dy, torch.bfloat16, "dy")
    _req(w, torch.bfloat16, "w")
    d = _cdesc(x_shape, w.shape, stride, pad)
    if tuple(dy.shape) != (d.N, d.P, d.Q, d.K):
        raise ValueError(f"dy shape {tuple(dy.shape)} != {(d.N, d.P, d.Q, d.K)}")
    dx = out if out is not None else torch.empty(tuple(x_shape), dtype=torch.bfloat16, device=dy.device)
    with _span("tc_conv_dgrad", _conv_flops(d), _conv_tag(d)):
        if relu_mask is not None:
            _req(relu_mask, torch.uint8, "relu_mask"); _contig(relu_mask, "relu_mask")
            if accumulate or tuple(relu_mask.shape) != (d.N, d.H, d.W, d.C // 8):
                raise ValueError("relu_mask: [N,H,W,C/8] of the input, no accumulation")
            check(lib.rnn_conv2d_dgrad_masked_bf16(d, ptr(dy), ptr(w), ptr(relu_mask), ptr(dx), cur_stream()),
                  "rnn_conv2d_dgrad_masked_bf16")
        else:
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
    with _span("tc_conv_wgrad", _conv_flops(d), _conv_tag(d)):
        check(lib.rnn_conv2d_wgrad_bf16(d, ptr(x), ptr(dy), ptr(dw), int(dw.dtype == torch.float32),
                                        int(accumulate), ptr(ws), nbytes, cur_stream()), "rnn_conv2d_wgrad_bf16")
    return dw


# =========================================================================================== helpers
def _is_f32(t: torch.Tensor) -> int:
    if t.dtype == torch.float32:
        return 1
    if t.dtype == torch.bfloat16:
        return 0
    raise TypeError(f"unsupported dtype {t.dtype}: kernels exist for float32 and bfloat16")


def _cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise _lib.RnnError("replicann_b200 kernels need CUDA tensors; there is no CPU path")


def _contig(t, name):
    if not t.is_contiguous():
        raise ValueError(f"{name} must be contiguous")


# =========================================================================================== fp32 path
def gemm_f32(a, b, *, a_mn=False, b_mn=False, bias=None, relu=False, out=None, accumulate=False):
    lib = _lib.load()
    _req(a, torch.float32, "a"); _req(b, torch.float32, "b")
    K, M = (a.shape if a_mn else a.shape[::-1])
    Kb, N = (b.shape if b_mn else b.shape[::-1])
    if K != Kb:
        raise ValueError(f"reduction mismatch {K} vs {Kb}")
    d = out if out is not None else torch.empty((M, N), dtype=torch.float32, device=a.device)
    # large enough and aligned: fp32-accurate GEMM on the tensor cores (three-piece bf16 split); else CUDA cores
    if (_f32_tc_enabled and not accumulate and M * N * K >= (1 << 22) and d.is_contiguous()
            and lib.rnn_gemm_f32_tc_supported(M, N, K, int(a_mn), int(b_mn))):
        if bias is not None:
            _req(bias, torch.float32, "bias")
        nbytes = lib.rnn_gemm_f32_tc_workspace(M, N, K, int(bias is not None or relu))
        ws = workspace(nbytes, a.device)
        with _span("tc_gemm", 12.0 * M * N * K, f"f32x3 M{M} N{N} K{K} a_mn{int(a_mn)} b_mn{int(b_mn)}"):
            check(lib.rnn_gemm_f32_tc(ptr(a), ptr(b), ptr(d), M, N, K, int(a_mn), int(b_mn), ptr(bias), int(relu),
                                      ptr(ws), nbytes, cur_stream()), "rnn_gemm_f32_tc")
        return d
    check(lib.rnn_gemm_f32(ptr(a), ptr(b), ptr(d), M, N, K, int(a_mn), int(b_mn), ptr(bias), int(relu),
                           int(accumulate), cur_stream()), "rnn_gemm_f32")
    return d


def conv2d_fprop_f32(x, w, bias=None, *, stride=1, pad=0):
    lib = _lib.load()
    _req(x, torch.float32, "x"); _req(w, torch.float32, "w")
    d = _cdesc(x.shape, w.shape, stride, pad)
    y = torch.empty((d.N, d.P, d.Q, d.K), dtype=torch.float32, device=x.device)
    check(lib.rnn_conv2d_fprop_f32(d, ptr(x), ptr(w), ptr(bias), ptr(y), cur_stream()), "rnn_conv2d_fprop_f32")
    return y


def conv2d_dgrad_f32(dy, w, x_shape, *, stride=1, pad=0):
    lib = _lib.load()
    _req(dy, torch.float32, "dy"); _req(w, torch.float32, "w")
    d = _cdesc(x_shape, w.shape, stride, pad)
    dx = torch.empty(tuple(x_shape), dtype=torch.float32, device=dy.device)
    check(lib.rnn_conv2d_dgrad_f32(d, ptr(dy), ptr(w), ptr(dx), cur_stream()), "rnn_conv2d_dgrad_f32")
    return dx


def conv2d_wgrad_f32(x, dy, w_shape, *, stride=1, pad=0, out=None, accumulate=False):
    lib = _lib.load()
    _req(x, torch.float32, "x"); _req(dy, torch.float32, "dy")
    d = _cdesc(x.shape, w_shape, stride, pad)
    dw = out if out is not None else torch.empty(tuple(w_shape), dtype=torch.float32, device=x.device)
    check(lib.rnn_conv2d_wgrad_f32(d, ptr(x), ptr(dy), ptr(dw), int(accumulate), cur_stream()), "rnn_conv2d_wgrad_f32")
    return dw


def colsum(x2d, *, out=None, accumulate=False):
    """out[c] (+)= sum_r x2d[r, c]; output dtype = input dtype."""
    lib = _lib.load()
    _cuda(x2d); _contig(x2d, "x")
    M, Cn = x2d.shape
    o = out if out is not None else torch.empty((Cn,), dtype=x2d.dtype, device=x2d.device)
    nbytes = lib.rnn_colsum_workspace(M, Cn)
    ws = workspace(nbytes, x2d.device)
    with _span('colsum', float(x2d.numel()) * x2d.element_size(), f"M{M} C{Cn}"):
        check(lib.rnn_colsum(ptr(x2d), M, Cn, ptr(o), int(accumulate), _is_f32(x2d), ptr(ws), nbytes, cur_stream()),
              "rnn_colsum")
    return o


# =========================================================================================== elementwise
def nchw_to_nhwc_bf16(x, cpad=None, mean=None, inv_std=None):
    lib = _lib.load()
    _req(x, torch.float32, "x")
    N, Cn, H, W = x.shape
    cp = cpad or Cn
    _norm_args(mean, inv_std, Cn)
    y = torch.empty((N, H, W, cp), dtype=torch.bfloat16, device=x.device)
    with _span("layout", 4.0 * N * Cn * H * W + 2.0 * N * H * W * cp, f"N{N} C{Cn} {H}x{W} -> C{cp}"):
        check(lib.rnn_nchw_f32_to_nhwc_bf16_norm(ptr(x), ptr(y), N, Cn, H, W, cp, ptr(mean), ptr(inv_std), cur_stream()),
              "rnn_nchw_f32_to_nhwc_bf16_norm")
    return y


def nhwc_bf16_to_nchw_f32(x, channels=None):
    lib = _lib.load()
    _req(x, torch.bfloat16, "x")
    N, H, W, cp = x.shape
    Cn = channels or cp
    y = torch.empty((N, Cn, H, W), dtype=torch.float32, device=x.device)
    check(lib.rnn_nhwc_bf16_to_nchw_f32(ptr(x), ptr(y), N, Cn, H, W, cp, cur_stream()), "rnn_nhwc_bf16_to_nchw_f32")
    return y


def relu_fwd(x):
    lib = _lib.load(); _cuda(x); _contig(x, "x")
    y = torch.empty_like(x)
    with _span('elementwise', float(2 * x.numel()) * x.element_size(), "relu_fwd"):
        check(lib.rnn_relu_fwd(ptr(x), ptr(y), x.numel(), _is_f32(x), cur_stream()), "rnn_relu_fwd")
    return y


def relu_bwd(dy, y):
    lib = _lib.load(); _cuda(dy, y)
    dx = torch.empty_like(y)
    with _span('elementwise', float(3 * y.numel()) * y.element_size(), "relu_bwd"):
        check(lib.rnn_relu_bwd(ptr(dy), ptr(y), ptr(dx), y.numel(), _is_f32(y), cur_stream()), "rnn_relu_bwd")
    return dx


def relu_bwd_colsum(dy, y, *, out, accumulate=False):
    """dx = dy * (y > 0) and out[c] (+)= column sums of dx over the last dimension, in one pass; returns dx.
    Falls back to the two separate kernels when the channel count is not a multiple of the vector width."""
    lib = _lib.load(); _cuda(dy, y)
    Cn = y.shape[-1]
    w = 4 if y.dtype == torch.float32 else 8
    if Cn % w or not dy.is_contiguous() or not y.is_contiguous() or out.dtype != y.dtype:
        dx = relu_bwd(dy, y)
        colsum(dx.reshape(-1, Cn), out=out, accumulate=accumulate)
        return dx
    M = y.numel() // Cn
    dx = torch.empty_like(y)
    nbytes = lib.rnn_colsum_workspace(M, Cn)
    ws = workspace(nbytes, y.device)
    with _span('colsum', float(3 * y.numel()) * y.element_size(), f"relu_bwd M{M} C{Cn}"):
        check(lib.rnn_relu_bwd_colsum(ptr(dy), ptr(y), ptr(dx), M, Cn, ptr(out), int(accumulate), _is_f32(y), ptr(ws),
                                      nbytes, cur_stream()), "rnn_relu_bwd_colsum")
    return dx


def tanh_fwd(x):
    lib = _lib.load(); _req(x, torch.float32, "x")
    y = torch.empty_like(x)
    check(lib.rnn_tanh_fwd_f32(ptr(x), ptr(y), x.numel(), cur_stream()), "rnn_tanh_fwd_f32")
    return y


def tanh_bwd(dy, y):
    lib = _lib.load(); _req(dy, torch.float32, "dy")
    dx = torch.empty_like(y)
    check(lib.rnn_tanh_bwd_f32(ptr(dy), ptr(y), ptr(dx), y.numel(), cur_stream()), "rnn_tanh_bwd_f32")
    return dx


def add(a, b, out=None):
    lib = _lib.load(); _req(a, torch.bfloat16, "a"); _req(b, torch.bfloat16, "b")
    y = out if out is not None else torch.empty_like(a)
    with _span('elementwise', float(6 * a.numel()), "add"):
        check(lib.rnn_add_bf16(ptr(a), ptr(b), ptr(y), a.numel(), cur_stream()), "rnn_add_bf16")
    return y


def cast_f32_to_bf16(x, out=None):
    lib = _lib.load(); _req(x, torch.float32, "x")
    y = out if out is not None else torch.empty(x.shape, dtype=torch.bfloat16, device=x.device)
    with _span('elementwise', float(6 * x.numel()), "cast"):
        check(lib.rnn_cast_f32_to_bf16(ptr(x), ptr(y), x.numel(), cur_stream()), "rnn_cast_f32_to_bf16")
    return y


def cast_bf16_to_f32(x, out=None):
    lib = _lib.load(); _req(x, torch.bfloat16, "x")
    y = out if out is not None else torch.empty(x.shape, dtype=torch.float32, device=x.device)
    with _span('elementwise', float(6 * x.numel()), "cast"):
        check(lib.rnn_cast_bf16_to_f32(ptr(x), ptr(y), x.numel(), cur_stream()), "rnn_cast_bf16_to_f32")
    return y


def dropout_fwd(x, p, seed, offset):
    lib = _lib.load(); _cuda(x); _contig(x, "x")
    y = torch.empty_like(x)
    mask = torch.empty(x.shape, dtype=torch.uint8, device=x.device)
    with _span('elementwise', float(x.numel()) * (2 * x.element_size() + 1), "dropout_fwd"):
        check(lib.rnn_dropout_fwd(ptr(x), ptr(y), ptr(mask), x.numel(), float(p), int(seed), int(offset), _is_f32(x),
                                  cur_stream()), "rnn_dropout_fwd")
    return y, mask


def dropout_bwd(dy, mask, p):
    lib = _lib.load(); _cuda(dy); _contig(dy, "dy")
    dx = torch.empty_like(dy)
    with _span('elementwise', float(dy.numel()) * (2 * dy.element_size() + 1), "dropout_bwd"):
        check(lib.rnn_dropout_bwd(ptr(dy), ptr(mask), ptr(dx), dy.numel(), float(p), _is_f32(dy), cur_stream()),
              "rnn_dropout_bwd")
    return dx


# =========================================================================================== normalisation
def bn_stats(x2d, gamma, beta, eps, momentum, running_mean, running_var):
    """x2d [M, C] bf16 -> (mean, invstd, scale, shift) fp32; updates running stats in place."""
    lib = _lib.load(); _req(x2d, torch.bfloat16, "x")
    M, Cn = x2d.shape
    st = torch.empty((4, Cn), dtype=torch.float32, device=x2d.device)
    nbytes = lib.rnn_bn_workspace(M, Cn)
    ws = workspace(nbytes, x2d.device)
    with _span("bn_stats", 2.0 * M * Cn, f"M{M} C{Cn}"):
      check(lib.rnn_bn_stats_bf16(ptr(x2d), M, Cn, ptr(gamma), ptr(beta), float(eps), float(momentum),
                                ptr(running_mean), ptr(running_var), ptr(st[0]), ptr(st[1]), ptr(st[2]), ptr(st[3]),
                                  ptr(ws), nbytes, cur_stream()), "rnn_bn_stats_bf16")
    return st[0], st[1], st[2], st[3]


def bn_eval_prepare(gamma, beta, running_mean, running_var, eps):
    lib = _lib.load()
    Cn = running_mean.numel()
    st = torch.empty((2, Cn), dtype=torch.float32, device=running_mean.device)
    check(lib.rnn_bn_eval_prepare(ptr(gamma), ptr(beta), ptr(running_mean), ptr(running_var), float(eps), Cn,
                                  ptr(st[0]), ptr(st[1]), cur_stream()), "rnn_bn_eval_prepare")
    return st[0], st[1]


def bn_apply(x2d, scale, shift, residual=None, relu=False, want_mask=False):
    """want_mask: also return the 1-bit-per-element ReLU mask (uint8 [M, C/8]); returns (y, mask) then."""
    lib = _lib.load(); _req(x2d, torch.bfloat16, "x")
    M, Cn = x2d.shape
    y = torch.empty_like(x2d)
    mask = torch.empty((M, Cn // 8), dtype=torch.uint8, device=x2d.device) if (want_mask and relu) else None
    with _span("bn_apply", 2.0 * M * Cn * (3 if residual is not None else 2), f"M{M} C{Cn} res{int(residual is not None)}"):
        check(lib.rnn_bn_apply_bf16(ptr(x2d), ptr(scale), ptr(shift), ptr(residual), int(relu), ptr(y), ptr(mask), M,
                                    Cn, cur_stream()), "rnn_bn_apply_bf16")
    return (y, mask) if want_mask else y


def bn_bwd(dy2d, x2d, y_out, mean, invstd, gamma, *, want_dres, dgamma, dbeta, accumulate, fwd_scale=None,
           fwd_shift=None, relu_mask=None):
    lib = _lib.load(); _req(dy2d, torch.bfloat16, "dy")
    M, Cn = x2d.shape
    dx = torch.empty_like(x2d)
    dres = torch.empty_like(x2d) if want_dres else None
    sums = torch.empty((5, Cn), dtype=torch.float32, device=x2d.device)
    nbytes = lib.rnn_bn_workspace(M, Cn)
    ws = workspace(nbytes, x2d.device)
    nread = 2 + (1 if y_out is not None else 0) + (0.0625 if relu_mask is not None else 0)
    with _span("bn_bwd", 2.0 * M * Cn * (2 * nread + 1 + (1 if want_dres else 0)), f"M{M} C{Cn} y{int(y_out is not None)} res{int(want_dres)}"):
        check(lib.rnn_bn_bwd_bf16(ptr(dy2d), ptr(x2d), ptr(y_out), ptr(mean), ptr(invstd), ptr(gamma), M, Cn,
                                  ptr(dx), ptr(dres), ptr(dgamma), ptr(dbeta),
                                  _is_f32(dgamma) if dgamma is not None else 1, int(accumulate), ptr(sums),
                                  ptr(fwd_scale), ptr(fwd_shift), ptr(relu_mask), ptr(ws), nbytes, cur_stream()),
              "rnn_bn_bwd_bf16")
    return dx, dres


def layernorm_fwd(x2d, residual, gamma, beta, eps, keep_sum=True):
    lib = _lib.load(); _cuda(x2d); _contig(x2d, "x")
    M, E = x2d.shape
    y = torch.empty_like(x2d)
    xsum = torch.empty_like(x2d) if (residual is not None and keep_sum) else None
    st = torch.empty((2, M), dtype=torch.float32, device=x2d.device)
    with _span('layernorm', float(x2d.numel()) * x2d.element_size() * (3 if residual is not None else 2), f"fwd M{M} E{E}"):
        check(lib.rnn_layernorm_fwd(ptr(x2d), ptr(residual), ptr(gamma), ptr(beta), float(eps), ptr(y), ptr(xsum),
                                    ptr(st[0]), ptr(st[1]), M, E, _is_f32(x2d), cur_stream()), "rnn_layernorm_fwd")
    return y, (xsum if xsum is not None else x2d), st[0], st[1]


def layernorm_bwd(dy2d, xin2d, gamma, mean, rstd, dgamma, dbeta, accumulate):
    lib = _lib.load(); _cuda(dy2d); _contig(dy2d, "dy")
    M, E = xin2d.shape
    dx = torch.empty_like(xin2d)
    nbytes = lib.rnn_layernorm_bwd_workspace(M, E)
    ws = workspace(nbytes, xin2d.device)
    with _span('layernorm', float(3 * xin2d.numel()) * xin2d.element_size(), f"bwd M{M} E{E}"):
        check(lib.rnn_layernorm_bwd(ptr(dy2d), ptr(xin2d), ptr(gamma), ptr(mean), ptr(rstd), ptr(dx), ptr(dgamma),
                                    ptr(dbeta), int(accumulate), M, E, _is_f32(xin2d), ptr(ws), nbytes, cur_stream()),
              "rnn_layernorm_bwd")
    return dx


# =========================================================================================== pooling
def _pool_args(x, k, s, p):
    N, H, W, Cn = x.shape
    kh, kw = (k, k) if isinstance(k, int) else k
    sh, sw = (s, s) if isinstance(s, int) else s
    ph, pw = (p, p) if isinstance(p, int) else p
    P = (H + 2 * ph - kh) // sh + 1
    Q = (W + 2 * pw - kw) // sw + 1
    return N, H, W, Cn, kh, kw, sh, sw, ph, pw, P, Q


def maxpool2d_fwd(x, k, s, p, return_indices=True, return_window_idx=False):
    """Returns (y, idx int32 or None, window_idx uint8 or None)."""
    lib = _lib.load(); _cuda(x); _contig(x, "x")
    N, H, W, Cn, kh, kw, sh, sw, ph, pw, P, Q = _pool_args(x, k, s, p)
    y = torch.empty((N, P, Q, Cn), dtype=x.dtype, device=x.device)
    idx = torch.empty((N, P, Q, Cn), dtype=torch.int32, device=x.device) if return_indices else None
    widx = torch.empty((N, P, Q, Cn), dtype=torch.uint8, device=x.device) if return_window_idx else None
    with _span('maxpool', float(x.numel() + 2 * y.numel()) * x.element_size(), f"fwd {tuple(x.shape)} k{kh} s{sh}"):
        check(lib.rnn_maxpool2d_fwd(ptr(x), ptr(y), ptr(idx), ptr(widx), N, H, W, Cn, kh, kw, sh, sw, ph, pw, _is_f32(x),
                                    cur_stream()), "rnn_maxpool2d_fwd")
    return y, idx, widx


def bn_relu_maxpool2d_fwd(x, scale, shift, k, s, p):
    """relu(x * scale + shift) pooled in one pass (x: pre-BatchNorm activation, NHWC); returns (y, window_idx)."""
    lib = _lib.load(); _cuda(x); _contig(x, "x")
    _req(scale, torch.float32, "scale"); _req(shift, torch.float32, "shift")
    N, H, W, Cn, kh, kw, sh, sw, ph, pw, P, Q = _pool_args(x, k, s, p)
    y = torch.empty((N, P, Q, Cn), dtype=x.dtype, device=x.device)
    widx = torch.empty((N, P, Q, Cn), dtype=torch.uint8, device=x.device)
    with _span('maxpool', float(x.numel() + y.numel()) * x.element_size() + y.numel(), f"bn+relu+fwd {tuple(x.shape)} k{kh} s{sh}"):
        check(lib.rnn_bn_relu_maxpool2d_fwd(ptr(x), ptr(scale), ptr(shift), ptr(y), None, ptr(widx), N, H, W, Cn, kh, kw,
                                            sh, sw, ph, pw, _is_f32(x), cur_stream()), "rnn_bn_relu_maxpool2d_fwd")
    return y, widx


def maxpool2d_bwd(dy, idx, x_shape, k, s, p, window_idx=None, relu_y=None):
    """relu_y: the pooled output when the pool's input came from a ReLU whose backward is applied here (dy is masked
    by relu_y > 0 inside the kernel)."""
    lib = _lib.load(); _cuda(dy); _contig(dy, "dy")
    N, H, W, Cn = x_shape
    kh, kw = (k, k) if isinstance(k, int) else k
    sh, sw = (s, s) if isinstance(s, int) else s
    ph, pw = (p, p) if isinstance(p, int) else p
    dx = torch.empty(tuple(x_shape), dtype=dy.dtype, device=dy.device)
    with _span('maxpool', float(2 * dy.numel() + dx.numel()) * dy.element_size(), f"bwd {tuple(x_shape)} k{kh} s{sh}"):
        if relu_y is not None:
            _contig(relu_y, "relu_y")
            if relu_y.shape != dy.shape or relu_y.dtype != dy.dtype:
                raise ValueError("relu_y: the pooled output, same shape and dtype as dy")
            check(lib.rnn_maxpool2d_relu_bwd(ptr(dy), ptr(relu_y), ptr(idx) if window_idx is None else None,
                                             ptr(window_idx), ptr(dx), N, H, W, Cn, kh, kw, sh, sw, ph, pw, _is_f32(dy),
                                             cur_stream()), "rnn_maxpool2d_relu_bwd")
        else:
            check(lib.rnn_maxpool2d_bwd(ptr(dy), ptr(idx) if window_idx is None else None, ptr(window_idx), ptr(dx), N, H,
                                        W, Cn, kh, kw, sh, sw, ph, pw, _is_f32(dy), cur_stream()), "rnn_maxpool2d_bwd")
    return dx


def avgpool2d_fwd(x, k, s, p):
    lib = _lib.load(); _cuda(x); _contig(x, "x")
    N, H, W, Cn, kh, kw, sh, sw, ph, pw, P, Q = _pool_args(x, k, s, p)
    y = torch.empty((N, P, Q, Cn), dtype=x.dtype, device=x.device)
    with _span('avgpool', float(x.numel() + y.numel()) * x.element_size(), "fwd"):
        check(lib.rnn_avgpool2d_fwd(ptr(x), ptr(y), N, H, W, Cn, kh, kw, sh, sw, ph, pw, _is_f32(x), cur_stream()),
              "rnn_avgpool2d_fwd")
    return y


def avgpool2d_bwd(dy, x_shape, k, s, p):
    lib = _lib.load(); _cuda(dy); _contig(dy, "dy")
    N, H, W, Cn = x_shape
    kh, kw = (k, k) if isinstance(k, int) else k
    sh, sw = (s, s) if isinstance(s, int) else s
    ph, pw = (p, p) if isinstance(p, int) else p
    dx = torch.empty(tuple(x_shape), dtype=dy.dtype, device=dy.device)
    with _span('avgpool', float(dy.numel() + dx.numel()) * dy.element_size(), "bwd"):
        check(lib.rnn_avgpool2d_bwd(ptr(dy), ptr(dx), N, H, W, Cn, kh, kw, sh, sw, ph, pw, _is_f32(dy), cur_stream()),
              "rnn_avgpool2d_bwd")
    return dx


def global_avgpool_fwd(x):
    lib = _lib.load(); _cuda(x); _contig(x, "x")
    N, H, W, Cn = x.shape
    y = torch.empty((N, 1, 1, Cn), dtype=x.dtype, device=x.device)
    with _span('avgpool', float(x.numel() + y.numel()) * x.element_size(), "global fwd"):
        check(lib.rnn_global_avgpool_fwd(ptr(x), ptr(y), N, H * W, Cn, _is_f32(x), cur_stream()), "rnn_global_avgpool_fwd")
    return y


def global_avgpool_bwd(dy, x_shape):
    lib = _lib.load(); _cuda(dy); _contig(dy, "dy")
    N, H, W, Cn = x_shape
    dx = torch.empty(tuple(x_shape), dtype=dy.dtype, device=dy.device)
    with _span('avgpool', float(dy.numel() + dx.numel()) * dy.element_size(), "global bwd"):
        check(lib.rnn_global_avgpool_bwd(ptr(dy), ptr(dx), N, H * W, Cn, _is_f32(dy), cur_stream()),
              "rnn_global_avgpool_bwd")
    return dx


# =========================================================================================== softmax / loss
def softmax_fwd(x2d):
    lib = _lib.load(); _cuda(x2d); _contig(x2d, "x")
    y = torch.empty_like(x2d)
    check(lib.rnn_softmax_fwd(ptr(x2d), ptr(y), x2d.shape[0], x2d.shape[1], _is_f32(x2d), cur_stream()),
          "rnn_softmax_fwd")
    return y


def softmax_bwd(dy2d, y2d):
    lib = _lib.load(); _cuda(dy2d); _contig(dy2d, "dy")
    dx = torch.empty_like(y2d)
    check(lib.rnn_softmax_bwd(ptr(dy2d), ptr(y2d), ptr(dx), y2d.shape[0], y2d.shape[1], _is_f32(y2d), cur_stream()),
          "rnn_softmax_bwd")
    return dx


def softmax_ce(logits, target, grad_scale=None, want_grad=True):
    """Returns (loss scalar fp32 tensor, dlogits or None).  grad_scale defaults to 1/rows (mean reduction)."""
    lib = _lib.load(); _cuda(logits, target); _contig(logits, "logits")
    if target.dtype != torch.int64:
        raise TypeError("target must be int64 class indices")
    rows, cols = logits.shape
    row_loss = torch.empty((rows + 1,), dtype=torch.float32, device=logits.device)
    dl = torch.empty_like(logits) if want_grad else None
    gs = (1.0 / rows) if grad_scale is None else grad_scale
    with _span('softmax_ce', float(logits.numel() * (3 if want_grad else 2)) * logits.element_size(), f"{rows}x{cols}"):
        check(lib.rnn_softmax_ce(ptr(logits), ptr(target), rows, cols, float(gs), ptr(row_loss), ptr(row_loss[rows:]),
                                 ptr(dl), _is_f32(logits), cur_stream()), "rnn_softmax_ce")
    return row_loss[rows], dl


def embedding_fwd(ids, tok, pos=None):
    """ids int64 [B, T] -> [B, T, E] = tok[ids] (+ pos[:T])."""
    lib = _lib.load(); _cuda(ids, tok, pos); _contig(ids, "ids"); _contig(tok, "tok")
    if ids.dtype != torch.int64:
        raise TypeError("embedding indices must be int64")
    B, T = ids.shape
    V, E = tok.shape
    if pos is not None:
        _contig(pos, "pos")
        if pos.dtype != tok.dtype or pos.shape[0] < T or pos.shape[1] != E:
            raise ValueError(f"positional table {tuple(pos.shape)} / {pos.dtype} does not cover T={T}, E={E} of the token table")
    out = torch.empty((B, T, E), dtype=tok.dtype, device=tok.device)
    with _span("embedding", float(B * T * E * tok.element_size() * (3 if pos is not None else 2)), f"fwd N{B * T} V{V} E{E}"):
        check(lib.rnn_embedding_fwd(ptr(ids), ptr(tok), ptr(pos), ptr(out), B * T, T, V, E, _is_f32(tok), cur_stream()),
              "rnn_embedding_fwd")
    return out


def embedding_bwd(ids, dy, dtok=None, dpos=None, *, accumulate_tok=False, accumulate_pos=False):
    """Scatter-add of dy [B, T, E] into dtok [V, E] and / or column sums over the batch into dpos[:T]."""
    lib = _lib.load(); _cuda(ids, dy, dtok, dpos); _contig(dy, "dy")
    B, T = ids.shape
    E = dy.shape[-1]
    V = dtok.shape[0] if dtok is not None else 1
    nbytes = lib.rnn_embedding_bwd_workspace(V) if dtok is not None else 0
    ws = workspace(nbytes, dy.device)
    with _span("embedding", float(B * T * E * dy.element_size() * 2), f"bwd N{B * T} V{V} E{E}"):
        check(lib.rnn_embedding_bwd(ptr(ids), ptr(dy), ptr(dtok), ptr(dpos), B * T, T, V, E, _is_f32(dy),
                                    int(accumulate_tok), int(accumulate_pos), ptr(ws), nbytes, cur_stream()),
              "rnn_embedding_bwd")


def argmax(x2d):
    lib = _lib.load(); _cuda(x2d); _contig(x2d, "x")
    out = torch.empty((x2d.shape[0],), dtype=torch.int64, device=x2d.device)
    check(lib.rnn_argmax(ptr(x2d), ptr(out), x2d.shape[0], x2d.shape[1], _is_f32(x2d), cur_stream()), "rnn_argmax")
    return out


# =========================================================================================== optimizers
def sgd_step(w, buf, grad, wcopy, *, lr, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False,
             first_step=False, grad_scale=1.0):
    lib = _lib.load(); _req(w, torch.float32, "w")
    with _span('optimizer', float(w.numel()) * (8 + (8 if buf is not None else 0) + grad.element_size() + (wcopy.element_size() if wcopy is not None else 0)), f"sgd n{w.numel()}"):
        check(lib.rnn_sgd_step(ptr(w), ptr(buf), ptr(grad), int(grad.dtype == torch.bfloat16), ptr(wcopy),
                               int(wcopy is not None and wcopy.dtype == torch.bfloat16), w.numel(), float(lr),
                               float(momentum), float(dampening), float(weight_decay), int(nesterov), int(first_step),
                               float(grad_scale), cur_stream()), "rnn_sgd_step")


def adamw_step(w, m, v, grad, wcopy, *, lr, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2, step=1, grad_scale=1.0,
               step_dev=None):
    """step_dev: int64 device tensor holding the step count (>= 1 when the kernel runs); overrides `step`."""
    lib = _lib.load(); _req(w, torch.float32, "w")
    common = (ptr(w), ptr(m), ptr(v), ptr(grad), int(grad.dtype == torch.bfloat16), ptr(wcopy),
              int(wcopy is not None and wcopy.dtype == torch.bfloat16), w.numel(), float(lr),
              float(betas[0]), float(betas[1]), float(eps), float(weight_decay))
    if step_dev is not None:
        _req(step_dev, torch.int64, "step_dev")
        with _span('optimizer', float(w.numel()) * (24 + grad.element_size() + (wcopy.element_size() if wcopy is not None else 0)), f"adamw n{w.numel()}"):
            check(lib.rnn_adamw_step_dev(*common, ptr(step_dev), float(grad_scale), cur_stream()), "rnn_adamw_step_dev")
    else:
        with _span('optimizer', float(w.numel()) * (24 + grad.element_size() + (wcopy.element_size() if wcopy is not None else 0)), f"adamw n{w.numel()}"):
            check(lib.rnn_adamw_step(*common, int(step), float(grad_scale), cur_stream()), "rnn_adamw_step")


def counter_add(counter: torch.Tensor, inc: int = 1) -> None:
    _req(counter, torch.int64, "counter")
    check(_lib.load().rnn_counter_add(ptr(counter), int(inc), cur_stream()), "rnn_counter_add")


# =========================================================================================== attention
def _attn_desc(B, H, T, Tk, hs, q, k, v, z, dz, scale, mask, p_drop, seed, offset):
    from ._lib import AttnDesc
    d = AttnDesc()
    d.B, d.H, d.T, d.Tk, d.hs = B, H, T, Tk, hs
    d.q_bs, d.q_rs = q.stride(0), q.stride(1)
    d.k_bs, d.k_rs = k.stride(0), k.stride(1)
    d.v_bs, d.v_rs = v.stride(0), v.stride(1)
    d.z_bs, d.z_rs = z.stride(0), z.stride(1)
    if dz is not None:
        d.dz_bs, d.dz_rs = dz.stride(0), dz.stride(1)
    d.scale = float(scale)
    d.p_drop, d.seed, d.offset = float(p_drop), int(seed), int(offset)
    if mask is None:
        d.mask_kind = 0
    else:
        d.mask_kind = 1 if mask.dtype in (torch.bool, torch.uint8) else 2
        if mask.stride(-1) != 1 and mask.shape[-1] != 1:
            raise ValueError("mask must have unit stride in its last dimension")
        if mask.dim() == 2:
            if tuple(mask.shape) != (T, Tk):
                raise ValueError(f"mask shape {tuple(mask.shape)} != ({T}, {Tk}); use broadcast_mask()")
            d.mask_bs, d.mask_rs = 0, mask.stride(0)
        else:
            if mask.shape[0] not in (1, B) or tuple(mask.shape[1:]) != (T, Tk):
                raise ValueError(f"mask shape {tuple(mask.shape)} != ({B}, {T}, {Tk}); use broadcast_mask()")
            d.mask_bs, d.mask_rs = (mask.stride(0) if mask.shape[0] > 1 else 0), mask.stride(1)
    return d


def broadcast_mask(mask, q, k, v, n_heads, hs):
    """Validate q / k / v against each other and bring `mask` to the [B, T, Tk] view the kernels index.

    The reference applies the mask with `w.masked_fill(mask, -inf)` / `w += mask` on w [B, T, Tk]
    (nn/attention.py:39-42), so anything that broadcasts TO that shape is legal -- [Tk], [1, Tk], [T, 1],
    [B, 1, Tk], ... -- and anything else raises torch's broadcast RuntimeError.  The broadcast is a stride-0
    view (the kernels take a batch and a row stride); only a mask whose last dimension is itself broadcast is
    materialised."""
    for t, n in ((q, "q"), (k, "k"), (v, "v")):
        if t.dim() != 3:
            raise ValueError(f"{n} must be [B, T, H*hs], got shape {tuple(t.shape)}")
    B, T, Tk = q.shape[0], q.shape[1], k.shape[1]
    width = n_heads * hs
    if q.shape[-1] != width or k.shape[-1] != width or v.shape[-1] != width:
        raise ValueError(f"q / k / v must have {n_heads} x {hs} = {width} features, got "
                         f"{q.shape[-1]}, {k.shape[-1]}, {v.shape[-1]}")
    if k.shape[0] != B or v.shape[0] != B or v.shape[1] != Tk:
        raise ValueError(f"k / v must be [{B}, Tk, {width}] with the same Tk, got {tuple(k.shape)} and {tuple(v.shape)}")
    if mask is None:
        return None
    if mask.dim() > 3:
        raise RuntimeError(f"mask of shape {tuple(mask.shape)} does not broadcast to the attention weights [{B}, {T}, {Tk}]")
    m = mask.expand(B, T, Tk)   # raises torch's RuntimeError when the shapes do not broadcast
    return m if m.stride(-1) == 1 or Tk == 1 else m.contiguous()


def _req_dtype(t, dtype, name):
    if t.dtype != dtype:
        raise TypeError(f"{name} must be {dtype}, got {t.dtype}")


def _check_qkv(t, name):
    if t.dim() != 3 or t.stride(2) != 1:
        raise ValueError(f"{name} must be [B, T, H*hs] with unit stride in the last dimension")


def attention_fwd(q, k, v, n_heads, hs, scale, mask=None, p_drop=0.0, seed=0, offset=0, need_lse=True):
    """q [B,T,H*hs], k/v [B,Tk,H*hs] (views allowed) -> z [B,T,H*hs] contiguous, lse [B,H,T] fp32."""
    lib = _lib.load(); _cuda(q, k, v, mask)
    for t, n in ((q, "q"), (k, "k"), (v, "v")):
        _check_qkv(t, n)
    B, T, _ = q.shape
    Tk = k.shape[1]
    z = torch.empty((B, T, n_heads * hs), dtype=q.dtype, device=q.device)
    lse = torch.empty((B, n_heads, T), dtype=torch.float32, device=q.device) if need_lse else None
    d = _attn_desc(B, n_heads, T, Tk, hs, q, k, v, z, None, scale, mask, p_drop, seed, offset)
    check(lib.rnn_attention_fwd(d, ptr(q), ptr(k), ptr(v), ptr(mask), ptr(z), ptr(lse), _is_f32(q), cur_stream()),
          "rnn_attention_fwd")
    return z, lse


_attn_tc_enabled = os.environ.get("RNN_ATTN_TC", "1") != "0"
_f32_tc_enabled = os.environ.get("RNN_F32_TC", "1") != "0"


def attention_tc_ok(q, k, v, n_heads, hs) -> bool:
    """Whether the tensor-core attention pipeline takes these operands (bf16, head size % 64, T % 8, aligned strides)."""
    if not _attn_tc_enabled or q.dtype != torch.bfloat16:
        return False
    B, T, _ = q.shape
    z_rs, z_bs = n_heads * hs, T * n_heads * hs
    return (hs % 64 == 0 and T % 8 == 0 and k.shape[1] % 8 == 0 and z_rs % 8 == 0 and z_bs % 8 == 0
            and all(t.stride(0) % 8 == 0 and t.stride(1) % 8 == 0 and t.data_ptr() % 16 == 0 for t in (q, k, v)))


def attention_tc_fwd(q, k, v, n_heads, hs, scale, mask=None, p_drop=0.0, seed=0, offset=0):
    """Tensor-core attention forward: returns z [B,T,H*hs] and the saved probabilities (opaque uint8 buffer)."""
    lib = _lib.load(); _cuda(q, k, v, mask)
    for t, n in ((q, "q"), (k, "k"), (v, "v")):
        _check_qkv(t, n)
        _req_dtype(t, torch.bfloat16, n)
    B, T, _ = q.shape
    Tk = k.shape[1]
    z = torch.empty((B, T, n_heads * hs), dtype=q.dtype, device=q.device)
    d = _attn_desc(B, n_heads, T, Tk, hs, q, k, v, z, None, scale, mask, p_drop, seed, offset)
    saved = torch.empty(lib.rnn_attention_tc_saved_bytes(d), dtype=torch.uint8, device=q.device)
    nbytes = lib.rnn_attention_tc_workspace(d)
    ws = workspace(nbytes, q.device)
    with _span("attention_tc", 4.0 * B * n_heads * T * Tk * hs, f"fwd B{B} H{n_heads} T{T} Tk{Tk} hs{hs}"):
        check(lib.rnn_attention_tc_fwd(d, ptr(q), ptr(k), ptr(v), ptr(mask), ptr(z), ptr(saved), ptr(ws), nbytes,
                                       cur_stream()), "rnn_attention_tc_fwd")
    return z, saved


def attention_tc_bwd(q, k, v, saved, dz, n_heads, hs, scale, dq, dk, dv, mask=None, p_drop=0.0, seed=0, offset=0):
    lib = _lib.load(); _cuda(q, k, v, dz)
    _check_qkv(dz, "dz")
    for a, b, n in ((dq, q, "dq"), (dk, k, "dk"), (dv, v, "dv")):
        if a.stride() != b.stride() or a.shape != b.shape:
            raise ValueError(f"{n} must have the shape and strides of its forward tensor")
    B, T, _ = q.shape
    Tk = k.shape[1]
    d = _attn_desc(B, n_heads, T, Tk, hs, q, k, v, dz, dz, scale, mask, p_drop, seed, offset)
    d.z_bs, d.z_rs = T * n_heads * hs, n_heads * hs
    nbytes = lib.rnn_attention_tc_workspace(d)
    ws = workspace(nbytes, q.device)
    with _span("attention_tc", 8.0 * B * n_heads * T * Tk * hs, f"bwd B{B} H{n_heads} T{T} Tk{Tk} hs{hs}"):
        check(lib.rnn_attention_tc_bwd(d, ptr(q), ptr(k), ptr(v), ptr(saved), ptr(dz), ptr(dq), ptr(dk), ptr(dv),
                                       ptr(ws), nbytes, cur_stream()), "rnn_attention_tc_bwd")


_attn_fused_enabled = os.environ.get("RNN_ATTN_FUSED", "1") != "0"


def attention_fused_ok(q, k, v, n_heads, hs) -> bool:
    """Whether the fused tcgen05 attention takes these operands: bf16, head size 64 or 128, T / Tk multiples of 8,
    16-byte aligned rows.  (RNN_ATTN_FUSED=0 routes to the unfused tensor-core pipeline instead: A/B timing.)"""
    if not (_attn_fused_enabled and _attn_tc_enabled) or q.dtype != torch.bfloat16 or hs not in (64, 128):
        return False
    B, T, _ = q.shape
    return (T % 8 == 0 and k.shape[1] % 8 == 0 and B * n_heads < 65536
            and all(t.stride(0) % 8 == 0 and t.stride(1) % 8 == 0 and t.data_ptr() % 16 == 0 for t in (q, k, v)))


def attention_fused_fwd(q, k, v, n_heads, hs, scale, mask=None, p_drop=0.0, seed=0, offset=0):
    """Fused tensor-core attention forward: z [B,T,H*hs] and the log-sum-exp [B,H,T] (all that backward needs)."""
    lib = _lib.load(); _cuda(q, k, v, mask)
    for t, n in ((q, "q"), (k, "k"), (v, "v")):
        _check_qkv(t, n)
        _req_dtype(t, torch.bfloat16, n)
    B, T, _ = q.shape
    Tk = k.shape[1]
    z = torch.empty((B, T, n_heads * hs), dtype=q.dtype, device=q.device)
    lse = torch.empty((B, n_heads, T), dtype=torch.float32, device=q.device)
    d = _attn_desc(B, n_heads, T, Tk, hs, q, k, v, z, None, scale, mask, p_drop, seed, offset)
    with _span("attention_fused", 4.0 * B * n_heads * T * Tk * hs, f"fwd B{B} H{n_heads} T{T} Tk{Tk} hs{hs}"):
        check(lib.rnn_attention_fused_fwd(d, ptr(q), ptr(k), ptr(v), ptr(mask), ptr(z), ptr(lse), cur_stream()),
              "rnn_attention_fused_fwd")
    return z, lse


def attention_fused_bwd(q, k, v, z, dz, lse, n_heads, hs, scale, dq, dk, dv, mask=None, p_drop=0.0, seed=0, offset=0):
    lib = _lib.load(); _cuda(q, k, v, z, dz)
    _check_qkv(dz, "dz")
    for a, b, n in ((dq, q, "dq"), (dk, k, "dk"), (dv, v, "dv")):
        if a.stride() != b.stride() or a.shape != b.shape:
            raise ValueError(f"{n} must have the shape and strides of its forward tensor")
    B, T, _ = q.shape
    Tk = k.shape[1]
    delta = torch.empty((B, n_heads, T), dtype=torch.float32, device=q.device)
    d = _attn_desc(B, n_heads, T, Tk, hs, q, k, v, z, dz, scale, mask, p_drop, seed, offset)
    with _span("attention_fused", 10.0 * B * n_heads * T * Tk * hs, f"bwd B{B} H{n_heads} T{T} Tk{Tk} hs{hs}"):
        check(lib.rnn_attention_fused_bwd(d, ptr(q), ptr(k), ptr(v), ptr(mask), ptr(z), ptr(dz), ptr(lse), ptr(delta),
                                          ptr(dq), ptr(dk), ptr(dv), cur_stream()), "rnn_attention_fused_bwd")


def attention_f32_tc_ok(q, k, v, n_heads, hs) -> bool:
    """Whether the fp32 tensor-core attention pipeline takes these operands (fp32, head size / T / Tk % 8, aligned
    strides, and enough work to pay for the operand splits)."""
    if not _attn_tc_enabled or not _f32_tc_enabled or q.dtype != torch.float32:
        return False
    B, T, _ = q.shape
    Tk = k.shape[1]
    if B * n_heads * T * Tk * hs < (1 << 22) or B * n_heads >= 65536:
        return False
    return (hs % 8 == 0 and hs <= 128 and T % 8 == 0 and Tk % 8 == 0
            and all(t.stride(0) % 4 == 0 and t.stride(1) % 4 == 0 and t.data_ptr() % 16 == 0 for t in (q, k, v)))


def attention_f32_tc_fwd(q, k, v, n_heads, hs, scale, mask=None, p_drop=0.0, seed=0, offset=0):
    lib = _lib.load(); _cuda(q, k, v, mask)
    for t, n in ((q, "q"), (k, "k"), (v, "v")):
        _check_qkv(t, n)
        _req_dtype(t, torch.float32, n)
    B, T, _ = q.shape
    Tk = k.shape[1]
    z = torch.empty((B, T, n_heads * hs), dtype=q.dtype, device=q.device)
    d = _attn_desc(B, n_heads, T, Tk, hs, q, k, v, z, None, scale, mask, p_drop, seed, offset)
    saved = torch.empty(lib.rnn_attention_f32_tc_saved_bytes(d), dtype=torch.uint8, device=q.device)
    nbytes = lib.rnn_attention_f32_tc_workspace(d)
    ws = workspace(nbytes, q.device)
    with _span("attention_tc", 24.0 * B * n_heads * T * Tk * hs, f"f32 fwd B{B} H{n_heads} T{T} Tk{Tk} hs{hs}"):
        check(lib.rnn_attention_f32_tc_fwd(d, ptr(q), ptr(k), ptr(v), ptr(mask), ptr(z), ptr(saved), ptr(ws), nbytes,
                                           cur_stream()), "rnn_attention_f32_tc_fwd")
    return z, saved


def attention_f32_tc_bwd(q, k, v, saved, dz, n_heads, hs, scale, dq, dk, dv, mask=None, p_drop=0.0, seed=0, offset=0):
    lib = _lib.load(); _cuda(q, k, v, dz)
    _check_qkv(dz, "dz")
    for a, b, n in ((dq, q, "dq"), (dk, k, "dk"), (dv, v, "dv")):
        if a.stride() != b.stride() or a.shape != b.shape:
            raise ValueError(f"{n} must have the shape and strides of its forward tensor")
    B, T, _ = q.shape
    Tk = k.shape[1]
    d = _attn_desc(B, n_heads, T, Tk, hs, q, k, v, dz, dz, scale, mask, p_drop, seed, offset)
    d.z_bs, d.z_rs = T * n_heads * hs, n_heads * hs
    nbytes = lib.rnn_attention_f32_tc_workspace(d)
    ws = workspace(nbytes, q.device)
    with _span("attention_tc", 48.0 * B * n_heads * T * Tk * hs, f"f32 bwd B{B} H{n_heads} T{T} Tk{Tk} hs{hs}"):
        check(lib.rnn_attention_f32_tc_bwd(d, ptr(q), ptr(k), ptr(v), ptr(saved), ptr(dz), ptr(dq), ptr(dk), ptr(dv),
                                           ptr(ws), nbytes, cur_stream()), "rnn_attention_f32_tc_bwd")


def attention_bwd(q, k, v, z, dz, lse, n_heads, hs, scale, dq, dk, dv, mask=None, p_drop=0.0, seed=0, offset=0):
    """Writes dq/dk/dv (pre-allocated, same strides as q/k/v)."""
    lib = _lib.load(); _cuda(q, k, v, z, dz)
    _check_qkv(dz, "dz")
    for a, b, n in ((dq, q, "dq"), (dk, k, "dk"), (dv, v, "dv")):
        if a.stride() != b.stride() or a.shape != b.shape:
            raise ValueError(f"{n} must have the shape and strides of its forward tensor")
    B, T, _ = q.shape
    Tk = k.shape[1]
    delta = torch.empty((B, n_heads, T), dtype=torch.float32, device=q.device)
    d = _attn_desc(B, n_heads, T, Tk, hs, q, k, v, z, dz, scale, mask, p_drop, seed, offset)
    check(lib.rnn_attention_bwd(d, ptr(q), ptr(k), ptr(v), ptr(mask), ptr(z), ptr(dz), ptr(lse), ptr(delta), ptr(dq),
                                ptr(dk), ptr(dv), _is_f32(q), cur_stream()), "rnn_attention_bwd")
