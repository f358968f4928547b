"""CPU: the C-ABI shared library builds for sm_100a, loads without a GPU, and exports every symbol that
include/rnn_b200.h declares; argument checking works before any CUDA call is made."""
import ctypes as C
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


@pytest.fixture(scope="module")
def lib():
    from replicann_b200 import build
    build.build()
    from replicann_b200 import _lib
    return _lib.load()


def test_every_declared_symbol_is_exported(lib):
    from replicann_b200 import _lib
    header = (ROOT / "include" / "rnn_b200.h").read_text()
    declared = set(re.findall(r"\b(rnn_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 45
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in rnn_b200.h but not exported by librnn_b200.so"


def test_header_cites_reference_or_torch_for_every_entry_point():
    header = (ROOT / "include" / "rnn_b200.h").read_text()
    assert "attention.py:" in header and "transformer.py:" in header
    assert 'extern "C"' in header


def test_version_and_error_string(lib):
    assert lib.rnn_version() >= 100
    assert isinstance(lib.rnn_last_error(), bytes)


def test_shape_errors_are_reported_without_touching_the_gpu(lib):
    from replicann_b200 import _lib
    # N not a multiple of 8 -> RNN_ERR_SHAPE before any CUDA call
    rc = lib.rnn_gemm_bf16(None, None, None, 16, 5, 16, 0, 0, None, 0, 0, None, 0, None)
    assert rc == -1 and "requirement failed" in _lib.last_error()
    d = _lib.conv_desc(1, 8, 8, 3, 8, 3, 3, 1, 1)  # C=3 is not a multiple of 64
    assert lib.rnn_conv2d_fprop_bf16(d, None, None, None, None, None, None, 0, None) == -1
    assert lib.rnn_conv2d_wgrad_workspace(d) == 0
    # zero-length elementwise calls are no-ops
    assert lib.rnn_relu_fwd(None, None, 0, 0, None) == 0
    assert lib.rnn_sgd_step(None, None, None, 1, None, 1, 0, 0.1, 0.9, 0.0, 0.0, 0, 1, 1.0, None) == 0
    # static-scratch entry point refuses inputs beyond its documented limit
    assert lib.rnn_bn_stats_bf16_static(None, 10, 8192, None, None, 1e-5, 0.1, *([None] * 6), None) == -3


def test_no_cpu_fallback_in_product_path():
    import torch
    from replicann_b200 import _lib
    from replicann_b200 import functional as RF
    with pytest.raises(_lib.RnnError):
        RF.relu_fwd(torch.zeros(8))
    with pytest.raises(_lib.RnnError):
        RF.gemm(torch.zeros(8, 8, dtype=torch.bfloat16), torch.zeros(8, 8, dtype=torch.bfloat16))
    # the package never imports the oracle
    for f in (ROOT / "replicann_b200").rglob("*.py"):
        assert "oracle" not in f.read_text(), f


def test_sass_contains_blackwell_tensor_core_and_tma_instructions():
    """tcgen05.mma -> UTCHMMA, tcgen05.ld -> LDTM, TMA tiled + im2col -> UTMALDG (B200_PROFILING.md)."""
    import shutil
    import subprocess
    if shutil.which("cuobjdump") is None:
        pytest.skip("cuobjdump not available")
    obj = ROOT / "build" / "obj" / "tc_gemm.o"
    sass = subprocess.run(["cuobjdump", "-sass", str(obj)], capture_output=True, text=True).stdout
    for mnemonic in ("UTCHMMA", "LDTM", "UTMALDG.2D", "UTMALDG.4D.IM2COL"):
        assert mnemonic in sass, mnemonic
    assert "HMMA." not in sass.replace("UTCHMMA", "")  # no legacy mma.sync path
