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
