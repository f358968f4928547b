"""The tensor-core kernels have two launch forms: single CTAs and CTA pairs (cta_group::2, a 256-row tile whose
B operand is split between the two CTAs).  The library picks per shape; RNN_TC_PAIR=2 / 0 force one form wherever
the shape allows, so the whole tensor-core parity file is re-run under each in a child process (the switch is read
once per process)."""
import os
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["2", "0"])
def test_tc_parity_under_forced_launch_form(mode):
    env = dict(os.environ, RNN_TC_PAIR=mode)
    r = subprocess.run([sys.executable, "-m", "pytest", str(ROOT / "tests" / "test_tc_kernels.py"), "-m", "gpu", "-q",
                        "-x", "--timeout", "600", "-p", "no:cacheprovider"],
                       cwd=ROOT, env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]


@pytest.mark.gpu
def test_conv_parity_with_the_filter_resident_form_forced():
    """Convolutions whose filter slice fits in shared memory keep it there for the life of the CTA and stream only A
    tiles (the library picks this when a CTA walks many tiles: the batch-256 shapes of tests/test_tc_real_shapes.py).
    RNN_TC_BRES=2 forces the form on the small parity cases too, RNN_TC_BRES=0 switches it off."""
    for mode in ("2", "0"):
        env = dict(os.environ, RNN_TC_BRES=mode)
        r = subprocess.run([sys.executable, "-m", "pytest", str(ROOT / "tests" / "test_tc_kernels.py"), "-m", "gpu", "-q",
                            "-x", "-k", "conv", "--timeout", "600", "-p", "no:cacheprovider"],
                           cwd=ROOT, env=env, capture_output=True, text=True, timeout=900)
        assert r.returncode == 0, f"RNN_TC_BRES={mode}\n" + r.stdout[-3000:] + r.stderr[-2000:]


@pytest.mark.gpu
def test_conv_parity_with_two_ctas_per_sm():
    """Two CTAs per SM for the N = 64 convolution tiles (each with a three-stage ring and one epilogue group): the
    default takes it for every such launch except the statistics epilogue; RNN_TC_DUAL=1 includes that one, =0 is the
    single-CTA form.  It only engages at large tile counts, so the batch-256 C64 3x3 and VGG 224^2 cases (fprop with
    the statistics epilogue, dgrad, wgrad) are the parity check, under both non-default settings."""
    for mode in ("1", "0"):
        env = dict(os.environ, RNN_TC_DUAL=mode)
        r = subprocess.run([sys.executable, "-m", "pytest", str(ROOT / "tests" / "test_tc_real_shapes.py"), "-m", "gpu",
                            "-q", "-x", "-k", "r50_s1_3x3_c64", "--timeout", "600", "-p", "no:cacheprovider"],
                           cwd=ROOT, env=env, capture_output=True, text=True, timeout=900)
        assert r.returncode == 0, f"RNN_TC_DUAL={mode}\n" + r.stdout[-3000:] + r.stderr[-2000:]
