"""ctypes binding of librnn_b200.so (the C ABI declared in include/rnn_b200.h).

There is deliberately no fallback: if the shared library is missing or a call fails, the caller gets an
exception.  torch is used only for device memory and streams.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_PKG = Path(__file__).resolve().parent
LIB_PATH = _PKG / "_lib" / "librnn_b200.so"

_lib = None


class RnnError(RuntimeError):
    pass


class ConvDesc(C.Structure):
    _fields_ = [(n, C.c_int) for n in
                ("N", "H", "W", "C", "K", "R", "S", "stride_h", "stride_w", "pad_h", "pad_w", "P", "Q")]


def conv_desc(N, H, W, Cin, K, R, S, stride, pad) -> ConvDesc:
    sh, sw = (stride, stride) if isinstance(stride, int) else stride
    ph, pw = (pad, pad) if isinstance(pad, int) else pad
    P = (H + 2 * ph - R) // sh + 1
    Q = (W + 2 * pw - S) // sw + 1
    return ConvDesc(N, H, W, Cin, K, R, S, sh, sw, ph, pw, P, Q)


_vp, _i, _f, _sz, _ll = C.c_void_p, C.c_int, C.c_float, C.c_size_t, C.c_longlong
_cdp = C.POINTER(ConvDesc)

# name -> (restype, argtypes).  Kept in one table so tests can check it against include/rnn_b200.h.
SIGNATURES = {
    "rnn_last_error": (C.c_char_p, []),
    "rnn_version": (_i, []),
    "rnn_sm_count": (_i, []),
    "rnn_gemm_bf16_workspace": (_sz, [_i, _i, _i, _i, _i]),
    "rnn_gemm_bf16": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _vp, _i, _i, _vp, _sz, _vp]),
    "rnn_conv2d_fprop_bf16": (_i, [_cdp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp]),
    "rnn_conv2d_dgrad_bf16": (_i, [_cdp, _vp, _vp, _vp, _i, _vp]),
    "rnn_conv2d_wgrad_workspace": (_sz, [_cdp]),
    "rnn_conv2d_wgrad_bf16": (_i, [_cdp, _vp, _vp, _vp, _i, _i, _vp, _sz, _vp]),
}


def load(path: os.PathLike | None = None):
    """Load the shared library (once) and attach argtypes.  Raises RnnError if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    p = Path(path) if path else LIB_PATH
    if not p.exists():
        raise RnnError(
            f"{p} not found: build it with `python -m replicann_b200.build` (nvcc, sm_100a). "
            "replicann_b200 has no CPU or PyTorch fallback path.")
    lib = C.CDLL(str(p))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def last_error() -> str:
    return load().rnn_last_error().decode()


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        raise RnnError(f"{what or 'rnn call'} failed with status {rc}: {last_error()}")


def ptr(t):
    """Device pointer of a torch tensor (or None)."""
    if t is None:
        return None
    return C.c_void_p(t.data_ptr())


def cur_stream():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)
