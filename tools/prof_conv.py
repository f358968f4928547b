"""Run one convolution shape a few times (for ncu / quick timing).
usage: python tools/prof_conv.py KIND N H W C K R stride pad [iters]     KIND in fprop|dgrad|wgrad"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from replicann_b200 import functional as RF  # noqa: E402

kind = sys.argv[1]
N, H, W, C, K, R, stride, pad = (int(v) for v in sys.argv[2:10])
iters = int(sys.argv[10]) if len(sys.argv) > 10 else 5
x = torch.randn(N, H, W, C, device="cuda").to(torch.bfloat16)
w = (torch.randn(K, R, R, C, device="cuda") * 0.05).to(torch.bfloat16)
P = (H + 2 * pad - R) // stride + 1
dy = torch.randn(N, P, P, K, device="cuda").to(torch.bfloat16)


def run():
    if kind == "fprop":
        return RF.conv2d_fprop(x, w, stride=stride, pad=pad)
    if kind == "dgrad":
        return RF.conv2d_dgrad(dy, w, tuple(x.shape), stride=stride, pad=pad)
    return RF.conv2d_wgrad(x, dy, tuple(w.shape), stride=stride, pad=pad)


for _ in range(3):
    run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(iters):
    run()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / iters
fl = 2.0 * N * P * P * K * C * R * R
print(f"{kind} N{N} {H}x{W} C{C} K{K} {R}x{R} s{stride} p{pad}: {ms*1e3:.1f} us  {fl/ms/1e9:.1f} TFLOP/s")
