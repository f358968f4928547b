"""Run N forward+backward passes of the reference's TransformerEncoder block (for ncu launch lists).
usage: python tools/run_encoder.py [S|L] [bf16|f32] [steps]"""
import sys
from pathlib import Path
import torch
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from replicann_b200.arch.transformer import TransformerEncoder  # noqa: E402

tag = sys.argv[1] if len(sys.argv) > 1 else "L"
dt = torch.float32 if (len(sys.argv) > 2 and sys.argv[2] == "f32") else torch.bfloat16
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 4
B, T, E, H = {"S": (64, 128, 512, 8), "L": (8, 512, 768, 12)}[tag]
enc = TransformerEncoder(H, E).cuda().to(dt).train()
x = torch.randn(B, T, E, device="cuda", dtype=dt, requires_grad=True)
for _ in range(steps):
    for p in enc.parameters():
        p.grad = None
    enc(x).float().sum().backward()
torch.cuda.synchronize()
print("ok", tag, dt, steps)
