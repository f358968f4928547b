"""Attention forward + backward at the survey's Track-R shapes S (B64 T128 E512 H8) and L (B8 T512 E768 H12), bf16:
the fused tcgen05 kernels (rnn_attention_fused_*) against the unfused tensor-core pipeline (rnn_attention_tc_*, six
batched GEMMs + row kernels with [B,H,T,T] tensors in HBM).  CUDA events, 3 warm-up + 20 timed, L2 flushed between
timed launches.   usage: python tools/bench_attention.py [out.tsv] [--once]   (--once: one pass of each, for ncu)"""
import math
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from replicann_b200 import functional as RF  # noqa: E402

once = "--once" in sys.argv
out = [a for a in sys.argv[1:] if not a.startswith("--")]
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timed(fn, iters=20):
    if once:
        fn()
        torch.cuda.synchronize()
        return 0.0
    for _ in range(3):
        fn()
    pairs = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        pairs.append((e0, e1))
    torch.cuda.synchronize()
    return sum(a.elapsed_time(b) for a, b in pairs) / iters


rows = ["shape\tpath\tfwd_ms\tbwd_ms\tfwd+bwd_ms\tTFLOP/s(fwd+bwd, 4+8 B*H*T*T*hs... see note)\tscore_tensor_MB"]
for tag, (B, T, E, H, p_drop, masked) in {"S": (64, 128, 512, 8, 0.0, False), "L": (8, 512, 768, 12, 0.0, False),
                                          "L+mask": (8, 512, 768, 12, 0.0, True), "L+dropout": (8, 512, 768, 12, 0.1, False),
                                          "L+mask+dropout": (8, 512, 768, 12, 0.1, True)}.items():
    hs = E // H
    g = torch.Generator().manual_seed(1)
    qkv = (torch.randn(B, T, 3 * E, generator=g) * 0.5).to(torch.bfloat16).cuda()
    q, k, v = qkv[..., :E], qkv[..., E:2 * E], qkv[..., 2 * E:]
    dz = (torch.randn(B, T, E, generator=g) * 0.1).to(torch.bfloat16).cuda()
    mask = torch.tril(torch.ones(T, T)).cuda() if masked else None
    scale = 1.0 / math.sqrt(E)
    dqkv = torch.empty_like(qkv)
    dq, dk, dv = dqkv[..., :E], dqkv[..., E:2 * E], dqkv[..., 2 * E:]
    flops = 12.0 * B * H * T * T * hs          # 2 products forward, 4 backward (algorithmic; recomputation not counted)
    z, lse = RF.attention_fused_fwd(q, k, v, H, hs, scale, mask, p_drop, 1, 1)
    f1 = timed(lambda: RF.attention_fused_fwd(q, k, v, H, hs, scale, mask, p_drop, 1, 1))
    b1 = timed(lambda: RF.attention_fused_bwd(q, k, v, z, dz, lse, H, hs, scale, dq, dk, dv, mask, p_drop, 1, 1))
    rows.append(f"{tag} B{B} T{T} H{H} hs{hs}\tfused\t{f1:.4f}\t{b1:.4f}\t{f1 + b1:.4f}\t"
                f"{flops / max(f1 + b1, 1e-9) / 1e9:.1f}\t0")
    z2, saved = RF.attention_tc_fwd(q, k, v, H, hs, scale, mask, p_drop, 1, 1)
    f2 = timed(lambda: RF.attention_tc_fwd(q, k, v, H, hs, scale, mask, p_drop, 1, 1))
    b2 = timed(lambda: RF.attention_tc_bwd(q, k, v, saved, dz, H, hs, scale, dq, dk, dv, mask, p_drop, 1, 1))
    rows.append(f"{tag} B{B} T{T} H{H} hs{hs}\tunfused\t{f2:.4f}\t{b2:.4f}\t{f2 + b2:.4f}\t"
                f"{flops / max(f2 + b2, 1e-9) / 1e9:.1f}\t{B * H * T * T * 2 / 1e6:.0f}")
text = "\n".join(rows) + "\n"
if out:
    Path(out[0]).write_text(text)
print(text)
