"""Where the time of a TransformerEncoder fwd+bwd goes (per-kernel-family CUDA-event spans), eager.
usage: python tools/prof_transformer.py [S|L] [bf16|f32]"""
import collections
import sys
from pathlib import Path
import torch
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from replicann_b200 import functional as RF  # noqa: E402
from replicann_b200.arch.transformer import TransformerEncoder  # noqa: E402

tag = sys.argv[1] if len(sys.argv) > 1 else "L"
dt = torch.float32 if (len(sys.argv) > 2 and sys.argv[2] == "f32") else torch.bfloat16
B, T, E, H = {"S": (64, 128, 512, 8), "L": (8, 512, 768, 12)}[tag]
enc = TransformerEncoder(H, E).cuda().to(dt).train()
x = torch.randn(B, T, E, device="cuda", dtype=dt, requires_grad=True)


def step():
    for p in enc.parameters():
        p.grad = None
    enc(x).float().sum().backward()


for _ in range(3):
    step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    step()
e1.record()
torch.cuda.synchronize()
print(f"encoder {tag} {dt}: {e0.elapsed_time(e1)/5:.3f} ms per fwd+bwd, launches/step {0}")
sink = []
RF.set_profile(sink)
n0 = RF.launch_count()
step()
torch.cuda.synchronize()
n1 = RF.launch_count()
RF.set_profile(None)
agg = collections.defaultdict(lambda: [0.0, 0])
for kind, work, a, b, t in sink:
    agg[kind][0] += a.elapsed_time(b); agg[kind][1] += 1
print(f"library launches per step: {n1 - n0}; instrumented spans: {len(sink)}")
for k, (ms, n) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    print(f"  {k:24s} {ms:8.3f} ms  {n:4d} calls")

# GPU-only time of the same step: replay it as a CUDA graph (no Python / launch overhead)
try:
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(2):
            step()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    gph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gph):
        step()
    for _ in range(3):
        gph.replay()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(10):
        gph.replay()
    e1.record()
    torch.cuda.synchronize()
    print(f"same step replayed as a CUDA graph: {e0.elapsed_time(e1)/10:.3f} ms")
except Exception as ex:  # noqa: BLE001
    print("graph capture failed:", repr(ex)[:300])

from torch.profiler import ProfilerActivity, profile  # noqa: E402
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    step()
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=25, max_name_column_width=60))
