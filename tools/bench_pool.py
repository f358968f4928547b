"""Times the ResNet stem pool kernels (3x3 / stride 2 / pad 1 over [256,112,112,64] bf16): forward, BatchNorm+ReLU
fused forward and backward, marching kernels against the one-row-per-CTA kernels (RNN_POOL_MARCH=0)."""
import os
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from replicann_b200 import functional as RF  # noqa: E402


def timed(fn, flush, n=20):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(n):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2] * 1e3


def main():
    N, H, W, C = 256, 112, 112, 64
    x = torch.randn(N, H, W, C, device="cuda").to(torch.bfloat16)
    scale = torch.rand(C, device="cuda") + 0.5
    shift = torch.randn(C, device="cuda") * 0.1
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    y, _, widx = RF.maxpool2d_fwd(x, 3, 2, 1, return_indices=False, return_window_idx=True)
    dy = torch.randn_like(y)
    xb, yb = x.numel() * 2, y.numel() * 2
    rows = []
    for mode in ("0", "1"):
        os.environ["RNN_POOL_MARCH"] = mode
        t_f = timed(lambda: RF.maxpool2d_fwd(x, 3, 2, 1, return_indices=False, return_window_idx=True), flush)
        t_b = timed(lambda: RF.bn_relu_maxpool2d_fwd(x, scale, shift, 3, 2, 1), flush)
        t_d = timed(lambda: RF.maxpool2d_bwd(dy, None, (N, H, W, C), 3, 2, 1, window_idx=widx), flush)
        byt = xb + yb + yb // 2
        rows.append(f"march={mode}\tfwd {t_f:.1f} us ({byt / t_f / 1e6:.2f} TB/s)\tbn+relu+fwd {t_b:.1f} us "
                    f"({byt / t_b / 1e6:.2f} TB/s)\tbwd {t_d:.1f} us ({byt / t_d / 1e6:.2f} TB/s)")
    print("\n".join(rows))
    if len(sys.argv) > 1:
        Path(sys.argv[1]).write_text("# stem pool [256,112,112,64] bf16, L2 flushed between launches, median of 20 "
                                     "(includes the output allocation)\n" + "\n".join(rows) + "\n")


if __name__ == "__main__":
    main()
