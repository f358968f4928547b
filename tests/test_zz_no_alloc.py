"""GPU test of rnn_bn_stats_bf16_static, the workspace-free BatchNorm-statistics entry point.

Some hosts of this library cannot allocate device memory on its behalf -- their HBM is fully owned by
another pool, or their binding can pass only the tensors themselves -- so they cannot supply the workspace
that rnn_bn_stats_bf16 asks for.  This test drives the static-scratch variant through the C ABI exactly as
such a host would (raw pointers, no workspace argument) and checks it against the regular entry point:
same mean / invstd / scale / shift and the same running-statistics update, on two back-to-back calls of
different shapes so that whatever the first call left in the shared scratch cannot leak into the second.
It also checks that an input wider than the built-in scratch supports is refused with a clear message.

Collected last (file name sorts after every other test file) so a run that stops at the first failure
has already executed everything else.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 1e-5  # both paths sum in fp32; only the partition of rows into blocks may differ


def _rel(a, b):
    return ((a - b).abs().max() / (b.abs().max() + 1e-12)).item()


def _run_both(lib, M, Cn, seed):
    from replicann_b200._lib import check
    g = torch.Generator().manual_seed(seed)
    x = (torch.randn(M, Cn, generator=g) * 1.7 + 0.3).to(torch.bfloat16).cuda()
    gamma = (torch.rand(Cn, generator=g) + 0.5).cuda()
    beta = torch.randn(Cn, generator=g).cuda()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    vp = lambda t: C.c_void_p(t.data_ptr())

    def outputs():
        return (torch.zeros(Cn, device="cuda"), torch.ones(Cn, device="cuda"),  # running mean / var
                *[torch.empty(Cn, device="cuda") for _ in range(4)])                 # mean invstd scale shift

    ref = outputs()
    nbytes = lib.rnn_bn_workspace(M, Cn)
    ws = torch.empty(max(nbytes, 1), dtype=torch.uint8, device="cuda")
    check(lib.rnn_bn_stats_bf16(vp(x), M, Cn, vp(gamma), vp(beta), 1e-5, 0.1, vp(ref[0]), vp(ref[1]), vp(ref[2]),
                                vp(ref[3]), vp(ref[4]), vp(ref[5]), vp(ws), nbytes, stream), "rnn_bn_stats_bf16")
    got = outputs()
    check(lib.rnn_bn_stats_bf16_static(vp(x), M, Cn, vp(gamma), vp(beta), 1e-5, 0.1, vp(got[0]), vp(got[1]),
                                       vp(got[2]), vp(got[3]), vp(got[4]), vp(got[5]), stream),
          "rnn_bn_stats_bf16_static")
    torch.cuda.synchronize()
    for name, a, b in zip(("running_mean", "running_var", "mean", "invstd", "scale", "shift"), got, ref):
        assert torch.isfinite(a).all(), name
        assert _rel(a, b) < TOL, (name, _rel(a, b))


def test_static_scratch_matches_workspace_entry_point_twice_in_a_row():
    from replicann_b200 import _lib
    lib = _lib.load()
    # call 1: wide and long (many partial rows -> the row-block count is capped by the built-in scratch)
    _run_both(lib, 50000, 2048, seed=1)
    # call 2, immediately after, different shape: must not see call 1's leftovers
    _run_both(lib, 777, 64, seed=2)


def test_static_scratch_at_its_widest_supported_input():
    """C = RNN_BN_STATIC_MAX_C exactly: the widest input the built-in scratch is sized for."""
    from replicann_b200 import _lib
    _run_both(_lib.load(), 3000, 4096, seed=3)


def test_static_scratch_on_a_second_device():
    """Every device holds its own copy of the scratch array and gets its own address lookup: the entry point must
    work with a device other than 0 current (hosts that drive several GPUs from one process)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from replicann_b200 import _lib
    with torch.cuda.device(1):
        torch.cuda.set_device(1)
        _run_both(_lib.load(), 4096, 256, seed=4)
    torch.cuda.set_device(0)


def test_static_scratch_rejects_inputs_beyond_its_limit():
    from replicann_b200 import _lib
    lib = _lib.load()
    Cn = 8192  # > RNN_BN_STATIC_MAX_C
    x = torch.zeros(16, Cn, dtype=torch.bfloat16, device="cuda")
    out = [torch.empty(Cn, device="cuda") for _ in range(6)]
    vp = lambda t: C.c_void_p(t.data_ptr())
    rc = lib.rnn_bn_stats_bf16_static(vp(x), 16, Cn, None, None, 1e-5, 0.1, vp(out[0]), vp(out[1]), vp(out[2]),
                                      vp(out[3]), vp(out[4]), vp(out[5]),
                                      C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == -3  # RNN_ERR_WORKSPACE
    assert "RNN_BN_STATIC_MAX_C" in _lib.last_error()
