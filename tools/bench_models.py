"""Step times of the other BASELINE.json configs and of the reference's transformer block (Track R):
LeNet-5 fp32 b64, AlexNet bf16 b256, VGG-16 bf16 b128, ResNet-50 bf16 b256, TransformerEncoder S / L.
usage: python tools/bench_models.py [out.tsv]   (one GPU; eager and CUDA-graph step where possible)"""
import sys
import time
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from replicann_b200.arch.alexnet import AlexNet  # noqa: E402
from replicann_b200.arch.lenet import LeNet5  # noqa: E402
from replicann_b200.arch.resnet import ResNet50  # noqa: E402
from replicann_b200.arch.transformer import TransformerEncoder  # noqa: E402
from replicann_b200.arch.vgg import VGG16  # noqa: E402
from replicann_b200.optim import SGD  # noqa: E402
from replicann_b200.training import Trainer  # noqa: E402


def timed(fn, warm=5, iters=10):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def cnn(name, model, x, y, graph, unit_count):
    opt = SGD(model.parameters(), lr=0.01, momentum=0.9)
    tr = Trainer(model, opt, cuda_graph=graph)
    ms = timed(lambda: tr.step(x, y))
    return f"{name}\t{'graph' if graph else 'eager'}\t{ms:.3f}\t{unit_count / ms * 1e3:.1f}\timages/s"


def main():
    out = open(sys.argv[1], "w") if len(sys.argv) > 1 else sys.stdout
    out.write("config\tmode\tms_per_step\trate\tunit\n")
    g = torch.Generator().manual_seed(1234)
    x = torch.randn(64, 1, 28, 28, generator=g).cuda().contiguous(memory_format=torch.channels_last)
    y = torch.randint(0, 10, (64,), generator=g).cuda()
    for graph in (False, True):
        out.write(cnn("LeNet-5 fp32 b64 1x28x28 (configs[0])", LeNet5().cuda(), x, y, graph, 64) + "\n")
    x = torch.randn(256, 3, 224, 224, generator=g).cuda()
    y = torch.randint(0, 1000, (256,), generator=g).cuda()
    for graph in (False, True):
        out.write(cnn("AlexNet bf16 fwd+bwd+SGD b256 (configs[1])", AlexNet(p_dropout=0.0).cuda().to(torch.bfloat16), x, y, graph, 256) + "\n")
    out.write(cnn("AlexNet bf16 fwd+bwd+SGD b256, dropout 0.5 (configs[1])", AlexNet().cuda().to(torch.bfloat16), x, y, False, 256) + "\n")
    for graph in (False, True):
        out.write(cnn("ResNet-50 bf16 b256 (configs[3])", ResNet50().cuda().to(torch.bfloat16), x, y, graph, 256) + "\n")
    x = x[:128].contiguous(); y = y[:128].contiguous()
    for graph in (False, True):
        out.write(cnn("VGG-16 bf16 b128 (configs[2])", VGG16(p_dropout=0.0).cuda().to(torch.bfloat16), x, y, graph, 128) + "\n")
    torch.cuda.empty_cache()
    for tag, (B, T, E, H) in {"S": (64, 128, 512, 8), "L": (8, 512, 768, 12)}.items():
        for dt in (torch.bfloat16, torch.float32):
            enc = TransformerEncoder(H, E).cuda().to(dt).train()
            xx = torch.randn(B, T, E, device="cuda", dtype=dt, requires_grad=True)
            from replicann_b200.optim import AdamW
            opt_ = AdamW(enc.parameters(), lr=1e-4)   # flat buckets: packed q/k/v projections, gradients in place

            def step(opt_=opt_, enc=enc, xx=xx):
                opt_.zero_grad()
                enc(xx).float().sum().backward()
            ms = timed(step, 3, 5)
            out.write(f"TransformerEncoder {tag} B{B} T{T} E{E} H{H} fwd+bwd {str(dt)[6:]}\teager\t{ms:.3f}\t{B * T / ms * 1e3:.0f}\ttokens/s\n")
            # the same step replayed as a CUDA graph: GPU time without the Python launch path
            gph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gph):
                step()
            ms = timed(gph.replay, 3, 10)
            out.write(f"TransformerEncoder {tag} B{B} T{T} E{E} H{H} fwd+bwd {str(dt)[6:]}\tgraph\t{ms:.3f}\t{B * T / ms * 1e3:.0f}\ttokens/s\n")
            del gph
    # the reference's decoder blocks (Track R rows a11 / a12), shape S, fp32, GPU time of forward + backward
    from replicann_b200.arch.transformer import TransformerCrossDecoder, TransformerDecoder
    B, T, E, H = 64, 128, 512, 8
    for name, cls in (("TransformerDecoder", TransformerDecoder), ("TransformerCrossDecoder", TransformerCrossDecoder)):
        # fp32: cast to bf16 the decoders' float tril mask raises the reference's TypeError (nn/attention.py:43-44)
        dec = cls(H, E, context_size=T).cuda().train()
        xx = torch.randn(B, T, E, device="cuda", requires_grad=True)
        kv = torch.randn(B, T, E, device="cuda")

        def dstep():
            for p in dec.parameters():
                p.grad = None
            y_ = dec(xx, kv, kv) if cls is TransformerCrossDecoder else dec(xx)
            y_.float().sum().backward()
        try:
            ms = timed(dstep, 3, 5)
            out.write(f"{name} S B{B} T{T} E{E} H{H} fwd+bwd float32\teager\t{ms:.3f}\t{B * T / ms * 1e3:.0f}\ttokens/s\n")
            gph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gph):
                dstep()
            ms = timed(gph.replay, 3, 10)
            out.write(f"{name} S B{B} T{T} E{E} H{H} fwd+bwd float32\tgraph\t{ms:.3f}\t{B * T / ms * 1e3:.0f}\ttokens/s\n")
            del gph
        except Exception as ex:  # noqa: BLE001  (e.g. the reference's bf16 TypeError for the float decoder mask)
            out.write(f"{name} S f32\t-\t-\t-\t{type(ex).__name__}: {str(ex)[:80]}\n")
    if out is not sys.stdout:
        out.close()


if __name__ == "__main__":
    main()
