"""Why do tensor-core kernels run slower inside a step than alone?  Time the C64 3x3 convolution (fprop) right after
different predecessor kernels, with CUDA events around the convolution only."""
import os
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from replicann_b200 import functional as RF  # noqa: E402

N, H, W, C = 256, 56, 56, 64
x = torch.randn(N, H, W, C, device="cuda").to(torch.bfloat16)
w = (torch.randn(64, 3, 3, C, device="cuda") * 0.05).to(torch.bfloat16)
big = torch.randn(N, H, W, 256, device="cuda").to(torch.bfloat16)           # 411 MB
scale = torch.rand(256, device="cuda") + 0.5
shift = torch.randn(256, device="cuda") * 0.1
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
xs = torch.randn(N, H, W, C, device="cuda").to(torch.bfloat16)
sc64, sh64 = torch.rand(64, device="cuda") + 0.5, torch.randn(64, device="cuda") * 0.1
w1 = (torch.randn(64, 1, 1, 256, device="cuda") * 0.05).to(torch.bfloat16)


def pre_none():
    pass


def pre_flush():
    flush.zero_()


def pre_bn_big():
    RF.bn_apply(big.reshape(-1, 256), scale, shift, None, True)


def pre_bn_x():      # the real predecessor: BatchNorm-apply that PRODUCES the convolution's input
    global x
    x = RF.bn_apply(xs.reshape(-1, C), sc64, sh64, None, True).reshape(N, H, W, C)


def pre_conv1x1():   # another tensor-core kernel
    RF.conv2d_fprop(big, w1)


huge = torch.empty(12 << 30, dtype=torch.uint8, device="cuda")            # 12 GB: a step's worth of other tensors


def pre_huge():      # touch 12 GB of other memory: TLB entries and L2 contents of x / w are gone
    huge.zero_()


def pre_huge_touch():   # the same, then one word per 2 MB page of x is read again (re-warms translations only)
    huge.zero_()
    xv = x.view(-1)
    xv[:: (1 << 20)].sum()


def run(pre, iters=20):
    ts = []
    for i in range(iters + 3):
        pre()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        RF.conv2d_fprop(x, w, stride=1, pad=1)
        e1.record()
        torch.cuda.synchronize()
        if i >= 3:
            ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


print(f"RNN_TC_DUAL={os.environ.get('RNN_TC_DUAL', '0')} RNN_TC_BRES={os.environ.get('RNN_TC_BRES', 'default')}")
for name, pre in (("nothing (L2 warm)", pre_none), ("256 MB memset (L2 flushed)", pre_flush),
                  ("bn_apply on a 411 MB tensor", pre_bn_big), ("bn_apply producing x", pre_bn_x),
                  ("1x1 convolution C256->64", pre_conv1x1), ("12 GB memset", pre_huge),
                  ("12 GB memset + page touch of x", pre_huge_touch)):
    print(f"  conv C64 3x3 after {name:32s}: {run(pre):7.1f} us")
