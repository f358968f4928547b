"""Experiment: is the N=64 tile bound by im2col TMA or by something else?  Conv 3x3 C64->K64 on 16 images
(L2-resident) vs the GEMM of the same shape (A materialised, 58 MB, L2-resident)."""
import sys
from pathlib import Path
import torch
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from replicann_b200 import functional as RF  # noqa: E402
from gemm_sweep import time_it  # noqa: E402

for nimg in (16, 32):
    x = torch.randn(nimg, 56, 56, 64, device="cuda").to(torch.bfloat16)
    w = (torch.randn(64, 3, 3, 64, device="cuda") * 0.05).to(torch.bfloat16)
    ms = time_it(lambda: RF.conv2d_fprop(x, w, stride=1, pad=1), iters=50)
    fl = 2.0 * nimg * 56 * 56 * 64 * 576
    print(f"conv3x3 C64 K64 N{nimg}: {ms*1e3:.1f} us {fl/ms/1e9:.0f} TF")
    M = nimg * 56 * 56
    a = torch.randn(M, 576, device="cuda").to(torch.bfloat16)
    b = torch.randn(64, 576, device="cuda").to(torch.bfloat16)
    ms = time_it(lambda: RF.gemm(a, b), iters=50)
    print(f"gemm M{M} N64 K576: {ms*1e3:.1f} us {fl/ms/1e9:.0f} TF")
    for K2 in (128, 256):
        w2 = (torch.randn(K2, 3, 3, 64, device="cuda") * 0.05).to(torch.bfloat16)
        ms = time_it(lambda: RF.conv2d_fprop(x, w2, stride=1, pad=1), iters=50)
        print(f"conv3x3 C64 K{K2} N{nimg}: {ms*1e3:.1f} us {fl*K2/64/ms/1e9:.0f} TF")
        b2 = torch.randn(K2, 576, device="cuda").to(torch.bfloat16)
        ms = time_it(lambda: RF.gemm(a, b2), iters=50)
        print(f"gemm M{M} N{K2} K576: {ms*1e3:.1f} us {fl*K2/64/ms/1e9:.0f} TF")
