"""Data-parallel step time of the other CNN configs (BASELINE.json configs[2]: VGG-16 bf16, 128 images/GPU).
Launch: python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \\
        tools/bench_dp.py vgg16 [steps]          (or plain `python tools/bench_dp.py vgg16` for one GPU)
Prints one line from rank 0: images/s over all ranks, device-timed, max over ranks."""
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from replicann_b200.optim import SGD  # noqa: E402
from replicann_b200.parallel import DataParallel  # noqa: E402
from replicann_b200.training import Trainer  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "vgg16"
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if name == "vgg16":
        from replicann_b200.arch.vgg import VGG16
        model, B = VGG16(p_dropout=0.5), 128
    elif name == "alexnet":
        from replicann_b200.arch.alexnet import AlexNet
        model, B = AlexNet(), 256
    else:
        raise SystemExit("model: vgg16 | alexnet")
    torch.manual_seed(0)
    model = model.cuda().to(torch.bfloat16).train()
    opt = SGD(model.parameters(), lr=0.01, momentum=0.9, weight_decay=5e-4)
    dp = DataParallel(model, opt) if world > 1 else None
    tr = Trainer(model, opt, data_parallel=dp, cuda_graph=True)
    g = torch.Generator().manual_seed(1234 + rank)
    x = torch.randn(B, 3, 224, 224, generator=g).cuda()
    y = torch.randint(0, 1000, (B,), generator=g).cuda()
    for _ in range(6):
        tr.step(x, y)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        tr.step(x, y)
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        ms = t.item() / steps
        print(f"{name}\tgpus {world}\tbatch/gpu {B}\tgraph {tr.graph_captured}\t{ms:.3f} ms/step\t"
              f"{B * world / ms * 1e3:.0f} images/s", flush=True)
    if world > 1:
        os._exit(0)


if __name__ == "__main__":
    main()
