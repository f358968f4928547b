"""cProfile of the eager Python launch path of a TransformerEncoder fwd+bwd (the GPU is not the bound there)."""
import cProfile
import pstats
import sys
from pathlib import Path
import torch
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from replicann_b200.arch.transformer import TransformerEncoder  # noqa: E402

B, T, E, H = 64, 128, 512, 8
enc = TransformerEncoder(H, E).cuda().to(torch.bfloat16).train()
x = torch.randn(B, T, E, device="cuda", dtype=torch.bfloat16, requires_grad=True)


def step():
    for p in enc.parameters():
        p.grad = None
    enc(x).float().sum().backward()


for _ in range(5):
    step()
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(50):
    step()
torch.cuda.synchronize()
pr.disable()
st = pstats.Stats(pr)
st.sort_stats("cumulative").print_stats(45)
