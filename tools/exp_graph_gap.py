"""How much does a kernel boundary cost inside a replayed CUDA graph?  Chain of N tiny dependent launches."""
import sys
from pathlib import Path
import torch
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from replicann_b200 import functional as RF  # noqa: E402

x = torch.randn(1024, device="cuda").to(torch.bfloat16)
y = torch.empty_like(x)
lib = RF._lib.load()
N = 500
s = torch.cuda.Stream()
with torch.cuda.stream(s):
    for _ in range(3):
        RF.check(lib.rnn_relu_fwd(RF.ptr(x), RF.ptr(y), x.numel(), 0, RF.cur_stream()))
torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g, stream=s):
    for i in range(N):
        a, b = (x, y) if i % 2 == 0 else (y, x)
        RF.check(lib.rnn_relu_fwd(RF.ptr(a), RF.ptr(b), x.numel(), 0, RF.cur_stream()))
for _ in range(3):
    g.replay()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    g.replay()
e1.record()
torch.cuda.synchronize()
print(f"{N} dependent tiny kernels in a graph: {e0.elapsed_time(e1)/10*1e3/N:.2f} us per kernel")
