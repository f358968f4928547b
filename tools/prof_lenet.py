"""Per-kernel time of one LeNet-5 fp32 b64 training step (torch profiler, eager)."""
import sys
from pathlib import Path
import torch
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from replicann_b200.arch.lenet import LeNet5  # noqa: E402
from replicann_b200.optim import SGD  # noqa: E402
from replicann_b200.training import Trainer  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402

g = torch.Generator().manual_seed(0)
x = torch.randn(64, 1, 28, 28, generator=g).cuda().contiguous(memory_format=torch.channels_last)
y = torch.randint(0, 10, (64,), generator=g).cuda()
m = LeNet5().cuda()
tr = Trainer(m, SGD(m.parameters(), lr=0.01, momentum=0.9), cuda_graph=False)
for _ in range(3):
    tr.step(x, y)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    tr.step(x, y)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=30, max_name_column_width=70))
