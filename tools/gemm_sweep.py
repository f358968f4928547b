"""GEMM microbench over operand layouts (K-major vs MN-major) and shapes: TFLOP/s of rnn_gemm_bf16.
usage: python tools/gemm_sweep.py"""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from replicann_b200 import functional as RF  # noqa: E402


def time_it(fn, iters=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    print("M\tN\tK\ta_mn\tb_mn\tus\tTFLOP/s")
    shapes = ((4096, 4096, 4096), (8192, 8192, 2048), (256, 2304, 50176), (2048, 2048, 12544), (16384, 256, 2304))
    if len(sys.argv) > 1 and sys.argv[1] == "conv":  # GEMMs with the shapes of ResNet convolutions (tiled TMA instead of im2col TMA)
        shapes = ((50176, 512, 1024), (802816, 64, 576), (200704, 128, 1152), (50176, 256, 2304), (12544, 512, 4608))
    for M, N, K in shapes:
        for a_mn, b_mn in ((0, 0), (0, 1), (1, 1)):
            a = torch.randn((K, M) if a_mn else (M, K), device="cuda").to(torch.bfloat16)
            b = torch.randn((K, N) if b_mn else (N, K), device="cuda").to(torch.bfloat16)
            ms = time_it(lambda: RF.gemm(a, b, a_mn=bool(a_mn), b_mn=bool(b_mn)))
            print(f"{M}\t{N}\t{K}\t{a_mn}\t{b_mn}\t{ms*1e3:.1f}\t{2.0*M*N*K/ms/1e9:.1f}", flush=True)


if __name__ == "__main__":
    main()
