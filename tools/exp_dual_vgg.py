"""VGG-16 conv1_1 -> conv1_2 forward through the modules, conv1_2 timed with events (RNN_TC_DUAL A/B)."""
import os
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from replicann_b200.arch.vgg import VGG16  # noqa: E402

m = VGG16(p_dropout=0.0).cuda().to(torch.bfloat16)
x = torch.randn(128, 3, 224, 224, device="cuda")
c1, c2 = m._features[0], m._features[1]
print("RNN_TC_DUAL =", os.environ.get("RNN_TC_DUAL", "default"), "emit", c1._emit_relu_mask, c2._emit_relu_mask, c2._grad_premasked)
for mode in ("no_grad", "grad"):
    ts = []
    for i in range(8):
        with torch.set_grad_enabled(mode == "grad"):
            y1 = c1(x)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            y2 = c2(y1)
            e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    print(f"  conv1_2 forward ({mode}): {ts[len(ts)//2]:8.1f} us")
