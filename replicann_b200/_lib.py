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


class AttnDesc(C.Structure):
    _fields_ = ([(n, C.c_int) for n in ("B", "H", "T", "Tk", "hs")] +
                [(n, C.c_longlong) for n in ("q_bs", "q_rs", "k_bs", "k_rs", "v_bs", "v_rs", "z_bs", "z_rs", "dz_bs", "dz_rs")] +
                [("scale", C.c_float), ("mask_kind", C.c_int), ("mask_bs", C.c_longlong), ("mask_rs", C.c_longlong),
                 ("p_drop", C.c_float), ("seed", C.c_ulonglong), ("offset", C.c_ulonglong)])


def conv_desc(N, H, W, Cin, K, R, S, stride, pad) -> ConvDesc:
    sh, sw = (stride, stride) if isinstance(stride, int) else stride
    ph, pw = (pad, pad) if isinstance(pad, int) else pad
    P = (H + 2 * ph - R) // sh + 1
    Q = (W + 2 * pw - S) // sw + 1
    return ConvDesc(N, H, W, Cin, K, R, S, sh, sw, ph, pw, P, Q)


HEADER_PATH = _PKG.parent / "include" / "rnn_b200.h"
_HEADER_FALLBACK = _PKG / "_lib" / "rnn_b200.h"  # copy placed next to the .so by the build


def _ctype(decl: str):
    d = " ".join(decl.replace("*", " * ").split())
    if "*" in d:
        if "rnn_conv_desc" in d:
            return C.POINTER(ConvDesc)
        if "rnn_attn_desc" in d:
            return C.POINTER(AttnDesc)
        if d.startswith("const char"):
            return C.c_char_p
        return C.c_void_p
    base = d.rsplit(" ", 1)[0] if " " in d else d  # drop the parameter name
    base = base.replace("const ", "").strip()
    table = {
        "cudaStream_t": C.c_void_p, "int": C.c_int, "float": C.c_float, "size_t": C.c_size_t,
        "long long": C.c_longlong, "unsigned long long": C.c_ulonglong, "void": None,
    }
    if base not in table:
        raise RnnError(f"rnn_b200.h: unknown parameter type in '{decl}'")
    return table[base]


def parse_header(path: os.PathLike | None = None) -> dict:
    """Parse include/rnn_b200.h into {name: (restype, [argtypes])}: the header is the single source of
    truth for the binding, so the ctypes signatures cannot drift from the C ABI."""
    import re
    hp = Path(path) if path else (HEADER_PATH if HEADER_PATH.exists() else _HEADER_FALLBACK)
    text = hp.read_text()
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    text = re.sub(r"//[^\n]*", " ", text)
    sigs = {}
    for m in re.finditer(r"(const char\s*\*|size_t|unsigned long long|int)\s+(rnn_\w+)\s*\(([^;{}]*?)\)\s*;", text, flags=re.S):
        ret, name, args = m.group(1), m.group(2), m.group(3).strip()
        restype = (C.c_char_p if "char" in ret else C.c_size_t if ret == "size_t"
                   else C.c_ulonglong if "long long" in ret else C.c_int)
        argtypes = [] if args in ("", "void") else [_ctype(a.strip()) for a in args.split(",")]
        sigs[name] = (restype, argtypes)
    return sigs


SIGNATURES = parse_header()


def load(path: os.PathLike | None = None):
    """Load the shared library (once) and attach argtypes.  Raises RnnError if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    # RNN_B200_LIB: load another build of the same ABI (A/B timing of two builds on one box)
    p = Path(path) if path else Path(os.environ.get("RNN_B200_LIB") or LIB_PATH)
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


_raw_stream = None


def cur_stream():
    """cudaStream_t of torch's current stream on the current device.  Called once per kernel launch, so it goes
    through torch's raw accessors (a few hundred ns) rather than building a torch.cuda.Stream object (~20 us)."""
    global _raw_stream
    import torch
    if _raw_stream is None:
        get_raw = getattr(torch._C, "_cuda_getCurrentRawStream", None)
        get_dev = getattr(torch._C, "_cuda_getDevice", None)
        if get_raw is not None and get_dev is not None:
            _raw_stream = lambda: get_raw(get_dev())                      # noqa: E731
        else:
            _raw_stream = lambda: torch.cuda.current_stream().cuda_stream  # noqa: E731
    return C.c_void_p(_raw_stream())
