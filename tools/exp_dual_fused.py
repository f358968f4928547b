"""Two CTAs per SM (RNN_TC_DUAL=1) on the VGG-16 conv1_2 shape: plain, bias+ReLU+mask (fprop) and masked dgrad epilogues."""
import os
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from replicann_b200 import functional as RF  # noqa: E402

N, H, W, C = 128, 224, 224, 64
x = torch.randn(N, H, W, C, device="cuda").to(torch.bfloat16)
w = (torch.randn(64, 3, 3, C, device="cuda") * 0.05).to(torch.bfloat16)
b = torch.randn(64, device="cuda")
dy = torch.randn(N, H, W, 64, device="cuda").to(torch.bfloat16)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
_, mask = RF.conv2d_fprop_relu_mask(x, w, stride=1, pad=1, shift=b)


def t(fn, iters=8):
    ts = []
    for i in range(iters + 2):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        if i >= 2:
            ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


if len(sys.argv) > 1 and sys.argv[1] == "relu":   # the real thing: x is a ReLU output (half zeros), small weights
    x = torch.relu(x)
    w = (w * 0.5).contiguous()
    dy = (dy * 1e-3).to(torch.bfloat16)
print("RNN_TC_DUAL =", os.environ.get("RNN_TC_DUAL", "default"), sys.argv[1:])
print(f"  fprop plain            {t(lambda: RF.conv2d_fprop(x, w, stride=1, pad=1)):8.1f} us")
print(f"  fprop bias+relu        {t(lambda: RF.conv2d_fprop(x, w, stride=1, pad=1, shift=b, relu=True)):8.1f} us")
print(f"  fprop bias+relu+mask   {t(lambda: RF.conv2d_fprop_relu_mask(x, w, stride=1, pad=1, shift=b)):8.1f} us")
print(f"  dgrad plain            {t(lambda: RF.conv2d_dgrad(dy, w, tuple(x.shape), stride=1, pad=1)):8.1f} us")
print(f"  dgrad masked           {t(lambda: RF.conv2d_dgrad(dy, w, tuple(x.shape), stride=1, pad=1, relu_mask=mask)):8.1f} us")


def t_queue(fn, pre, iters=10):
    """The same, but everything enqueued back to back behind a spin kernel (the host far ahead of the device, no idle
    gaps), as inside a training step."""
    torch.cuda._sleep(40_000_000)
    evs = []
    for i in range(iters):
        pre()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        evs.append((e0, e1))
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) * 1e3 for a, b in evs)
    return ts[len(ts) // 2]


first = torch.randn(N, 3, H, W, device="cuda")
print(f"  queued, after flush:            fprop bias+relu+mask {t_queue(lambda: RF.conv2d_fprop_relu_mask(x, w, stride=1, pad=1, shift=b), lambda: flush.zero_()):8.1f} us")
print(f"  queued, after another tc conv:  fprop bias+relu+mask {t_queue(lambda: RF.conv2d_fprop_relu_mask(x, w, stride=1, pad=1, shift=b), lambda: RF.conv2d_fprop(x, w, stride=1, pad=1)):8.1f} us")
print(f"  queued, no predecessor:         fprop bias+relu+mask {t_queue(lambda: RF.conv2d_fprop_relu_mask(x, w, stride=1, pad=1, shift=b), lambda: None):8.1f} us")
