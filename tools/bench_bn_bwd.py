"""BatchNorm backward (reduce / finalize / apply) at the ResNet-50 tensor shapes (batch 256), with per-kernel device
times from the library's own event pairs.  Each timed call runs right after a dgrad-sized write of dy so that dy is
as L2-resident as it is inside a training step, everything else flushed.  RNN_BN_REDUCE_CTAS / RNN_BN_APPLY_CTAS
sweep the grid sizes.   usage: python tools/bench_bn_bwd.py [out.tsv]"""
import os
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from replicann_b200 import functional as RF  # noqa: E402

SHAPES = [(12544, 512), (50176, 256), (50176, 512), (12544, 2048), (200704, 128), (200704, 256), (50176, 1024),
          (802816, 64), (200704, 512), (802816, 256)]
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
rows = ["M\tC\tMB\tus\tGB/s(5 tensor passes)\tper-kernel us (warm, back to back)"]
for M, C in SHAPES:
    g = torch.Generator().manual_seed(0)
    x = torch.randn(M, C, generator=g).to(torch.bfloat16).cuda()
    dy_src = torch.randn(M, C, generator=g).to(torch.bfloat16).cuda()
    dy = torch.empty_like(dy_src)
    gamma = torch.ones(C, device="cuda")
    mean, invstd, scale, shift = RF.bn_stats(x, gamma, torch.zeros(C, device="cuda"), 1e-5, 0.1, None, None)
    dg, db = torch.empty(C, device="cuda"), torch.empty(C, device="cuda")

    def run():
        return RF.bn_bwd(dy, x, None, mean, invstd, gamma, want_dres=False, dgamma=dg, dbeta=db, accumulate=False,
                         fwd_scale=scale, fwd_shift=shift)
    for _ in range(3):
        run()
    pairs = []
    for _ in range(20):
        flush.zero_()
        dy.copy_(dy_src)     # dy freshly written, as after the dgrad kernel that produces it
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); run(); e1.record()
        pairs.append((e0, e1))
    torch.cuda.synchronize()
    us = sum(a.elapsed_time(b) for a, b in pairs) / len(pairs) * 1e3
    # per-kernel device times: 20 calls queued behind a spin so that the host is ahead of the device
    RF.timing_enable(True)
    torch.cuda._sleep(40_000_000)
    for _ in range(20):
        run()
    torch.cuda.synchronize()
    parts = " ".join(f"{k.replace('bn_bwd_', '')}={RF.timing_read(k)[0] / 20 * 1e3:.1f}"
                     for k in ("bn_bwd_reduce", "bn_bwd_finalize", "bn_bwd_apply") if RF.timing_read(k)[1])
    RF.timing_enable(False)
    rows.append(f"{M}\t{C}\t{M * C * 2 / 1e6:.1f}\t{us:.1f}\t{5 * M * C * 2 / us / 1e3:.0f}\t{parts}")
text = (f"# RNN_BN_REDUCE_CTAS={os.environ.get('RNN_BN_REDUCE_CTAS', 'default')} "
        f"RNN_BN_APPLY_CTAS={os.environ.get('RNN_BN_APPLY_CTAS', 'default')}\n" + "\n".join(rows) + "\n")
if len(sys.argv) > 1:
    Path(sys.argv[1]).write_text(text)
print(text)
