"""Per-kernel-family time of one VGG-16 bf16 b128 training step (eager, CUDA-event spans + torch profiler)."""
import sys
from pathlib import Path
import torch
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from replicann_b200.arch.alexnet import AlexNet  # noqa: E402
from replicann_b200.optim import SGD  # noqa: E402
from replicann_b200.training import Trainer  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402

g = torch.Generator().manual_seed(0)
x = torch.randn(256, 3, 224, 224, generator=g).cuda()
y = torch.randint(0, 1000, (256,), generator=g).cuda()
m = AlexNet(p_dropout=0.0).cuda().to(torch.bfloat16)
tr = Trainer(m, SGD(m.parameters(), lr=0.01, momentum=0.9), cuda_graph=False)
for _ in range(3):
    tr.step(x, y)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    tr.step(x, y)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=28, max_name_column_width=70))
