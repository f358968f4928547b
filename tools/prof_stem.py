"""Time the ResNet stem's memory-bound kernels alone: layout conversion, max-pool forward / backward.
usage: python tools/prof_stem.py"""
import sys
from pathlib import Path
import torch
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tools"))
from replicann_b200 import functional as RF  # noqa: E402
from replicann_b200 import _lib  # noqa: E402
from gemm_sweep import time_it  # noqa: E402

N = 256
x = torch.randn(N, 112, 112, 64, device="cuda").to(torch.bfloat16)
y, idx, widx = RF.maxpool2d_fwd(x, 3, 2, 1, return_indices=False, return_window_idx=True)
ms = time_it(lambda: RF.maxpool2d_fwd(x, 3, 2, 1, return_indices=False, return_window_idx=True), iters=20)
print(f"maxpool fwd  {ms*1e3:7.1f} us  {(x.numel()*2 + y.numel()*3)/ms/1e6:7.0f} GB/s")
dy = torch.randn_like(y)
ms = time_it(lambda: RF.maxpool2d_bwd(dy, None, tuple(x.shape), 3, 2, 1, window_idx=widx), iters=20)
print(f"maxpool bwd  {ms*1e3:7.1f} us  {(x.numel()*2 + y.numel()*3)/ms/1e6:7.0f} GB/s")
img = torch.randn(N, 3, 224, 224, device="cuda")
d = _lib.conv_desc(N, 224, 224, 3, 64, 7, 7, 2, 3)
if d is not None:
    out = RF.pack_first_input(img, d)
    ms = time_it(lambda: RF.pack_first_input(img, d), iters=20)
    print(f"pack input   {ms*1e3:7.1f} us  {(img.numel()*4 + out.numel()*2)/ms/1e6:7.0f} GB/s")
