"""Run the BatchNorm kernels on one activation shape a few times (for ncu / quick timing).
usage: python tools/prof_bn.py M C [res]"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from replicann_b200 import functional as RF  # noqa: E402

M, C = int(sys.argv[1]), int(sys.argv[2])
res = len(sys.argv) > 3
x = torch.randn(M, C, device="cuda").to(torch.bfloat16)
dy = torch.randn(M, C, device="cuda").to(torch.bfloat16)
r = torch.randn(M, C, device="cuda").to(torch.bfloat16) if res else None
g, b = torch.ones(C, device="cuda"), torch.zeros(C, device="cuda")
rm, rv = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
dg, db = torch.empty(C, device="cuda"), torch.empty(C, device="cuda")


def run():
    mean, invstd, scale, shift = RF.bn_stats(x, g, b, 1e-5, 0.1, rm, rv)
    y, bits = RF.bn_apply(x, scale, shift, r, True, want_mask=True)
    RF.bn_bwd(dy, x, None, mean, invstd, g, want_dres=res, dgamma=dg, dbeta=db, accumulate=False,
              fwd_scale=None if res else scale, fwd_shift=None if res else shift, relu_mask=bits if res else None)


for _ in range(3):
    run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    run()
e1.record()
torch.cuda.synchronize()
print(f"bn fwd+bwd M{M} C{C} res{int(res)}: {e0.elapsed_time(e1)/5*1e3:.1f} us")
