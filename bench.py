#!/usr/bin/env python
"""bench.py -- BASELINE.json's headline metric on B200: train images/sec, ResNet-50 bf16, batch 256/GPU.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K --warmup W   # CPU reference arm (torch-CPU model)
    python bench.py --config {resnet50,vgg16,alexnet,lenet5,conv_sweep} ...   # the other BASELINE.json configs

One "step" = one full training step (forward, softmax-CE, backward, gradient all-reduce when N > 1, fused
SGD-momentum update) of ResNet-50 on one synthetic batch.  Prints ONE JSON line on rank 0.  See DESIGN.md
section "Measurement" for how every field is produced.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

UNIT = "images/s"

# BASELINE.json `configs`.  The default (and the configuration the metric is quoted on) is configs[3]; the others are
# selectable so that the driver can run them with the same JSON contract.
CONFIGS = {
    "resnet50": {"metric": "train images/sec, ResNet-50 bf16", "batch": 256, "shape": (3, 224, 224), "classes": 1000,
                 "dtype": "bf16", "cpu_batch": 64, "cpu_steps": 8,
                 "workload": "ResNet-50 bf16 data-parallel training, batch 256/GPU, fused BN+ReLU, synthetic 3x224x224 (configs[3])"},
    "vgg16": {"metric": "train images/sec, VGG-16 bf16", "batch": 128, "shape": (3, 224, 224), "classes": 1000,
              "dtype": "bf16", "cpu_batch": 16, "cpu_steps": 8,
              "workload": "VGG-16 bf16 data-parallel training, batch 128/GPU, dropout 0.5, synthetic 3x224x224 (configs[2])"},
    "alexnet": {"metric": "train images/sec, AlexNet bf16", "batch": 256, "shape": (3, 224, 224), "classes": 1000,
                "dtype": "bf16", "cpu_batch": 256, "cpu_steps": 8,
                "workload": "AlexNet bf16 fwd+bwd (+ fused SGD), batch 256, dropout 0.5, synthetic 3x224x224 (configs[1])"},
    "lenet5": {"metric": "train images/sec, LeNet-5 fp32", "batch": 64, "shape": (1, 28, 28), "classes": 10,
               "dtype": "f32", "cpu_batch": 64, "cpu_steps": 3000,
               "workload": "LeNet-5 fp32 train step, batch 64, synthetic 1x28x28 (configs[0])"},
}


def _build(config: str, torch_ref: bool):
    """The network of a config: this repository's (CUDA) model, or its torch.nn CPU twin from oracle/."""
    if torch_ref:
        from oracle import torch_models as TM
        return {"resnet50": lambda: TM.RefResNet((3, 4, 6, 3), True, 1000),
                "vgg16": lambda: TM.RefVGG((64, 64, "M", 128, 128, "M", 256, 256, 256, "M", 512, 512, 512, "M",
                                            512, 512, 512, "M"), 1000),
                "alexnet": lambda: TM.RefAlexNet(1000), "lenet5": lambda: TM.RefLeNet5()}[config]()
    from replicann_b200.arch.alexnet import AlexNet
    from replicann_b200.arch.lenet import LeNet5
    from replicann_b200.arch.resnet import ResNet50
    from replicann_b200.arch.vgg import VGG16
    return {"resnet50": ResNet50, "vgg16": VGG16, "alexnet": AlexNet, "lenet5": LeNet5}[config]()


def _peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return {"hbm_gbs": d.get("hbm_gbs", 6650.0), "bf16_tflops": d.get("bf16_tflops", 1590.0),
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", 1400.0), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


# ------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons every 50 ms while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.idx}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1]))
            except ValueError:
                continue
            for n, v in zip(names, parts[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------ CPU arms
def _torch_cpu_steps(config: str, batch: int, steps: int, warmup: int, threads: int):
    """the Track-N CPU baseline is the torch-CPU model (oracle/torch_models.py -- torch is the
    library the reference's own arithmetic lives in) doing the same training step, fp32, with all host threads:
    forward, cross-entropy, backward, SGD-momentum.  Returns per-step seconds of the timed steps."""
    import torch
    import torch.nn.functional as F
    torch.set_num_threads(threads)
    cfg = CONFIGS[config]
    torch.manual_seed(0)
    net = _build(config, torch_ref=True).train()
    opt = torch.optim.SGD(net.parameters(), lr=0.01, momentum=0.9, weight_decay=1e-4)
    g = torch.Generator().manual_seed(1234)
    x = torch.randn(batch, *cfg["shape"], generator=g)
    y = torch.randint(0, cfg["classes"], (batch,), generator=g)
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        opt.zero_grad(set_to_none=True)
        loss = F.cross_entropy(net(x), y)
        loss.backward()
        opt.step()
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    return times


def _cpu_baseline(config: str, batch: int, steps: int, warmup: int) -> dict:
    cores = os.cpu_count() or 1
    ts = _torch_cpu_steps(config, batch, steps, warmup, cores)
    med, best = statistics.median(ts), min(ts)
    return {"value": batch / med, "unit": UNIT, "cores": cores, "kind": "port",
            "best": batch / best, "median_s_per_step": med, "best_s_per_step": best,
            "sample": f"{warmup} warm-up + {steps} timed {config} fp32 training steps (fwd, CE, bwd, SGD-momentum) of "
                      f"{batch} images on torch {__import__('torch').__version__} CPU, torch.set_num_threads({cores}); "
                      f"value = batch / median step"}


def run_reference(args) -> None:
    """--impl reference: the path's CPU implementation with all host threads, a bounded sample per step.
    The reference repository has no CNN code; its arithmetic library is torch, so the timed CPU
    implementation is the torch.nn model of oracle/torch_models.py (kind "port": our wiring, the reference's
    dependency doing every FLOP)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfg = CONFIGS[args.config]
    # exactly the K / W the caller asked for (the driver compares them with its own arm's); SURVEY 8d's ">= 5 timed,
    # 2 warm-up" is what the defaults and the driver's values give
    steps, warmup = max(args.steps, 1), max(args.warmup, 0)
    # one probe step sizes the per-step sample so that the whole run stays within ~3 minutes
    batch = cfg["cpu_batch"]
    probe = _torch_cpu_steps(args.config, batch, 1, 0, os.cpu_count() or 1)[0]
    while batch > 1 and probe * (steps + warmup) > 170:
        batch //= 2
        probe /= 2
    cb = _cpu_baseline(args.config, batch, steps, warmup)
    line = {
        "impl": "reference", "metric": cfg["metric"], "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": cb["median_s_per_step"] * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": cfg["workload"],
                   "per_step_sample": f"{batch} images (bounded CPU sample of the {cfg['batch']}-image batch)"},
        "cpu_baseline": cb,
        "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    _emit(line)


# ------------------------------------------------------------------------------------------ GPU arm
def run_gpu(args) -> None:
    import torch
    import torch.distributed as dist

    from replicann_b200 import functional as RF
    from replicann_b200.optim import SGD
    from replicann_b200.parallel import DataParallel
    from replicann_b200.training import Trainer

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- replicann_b200 has no CPU path (use --impl reference for the CPU arm)")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    if args.gpus != world:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {args.gpus}")

    cfg = CONFIGS[args.config]
    B = args.batch or cfg["batch"]
    bf16 = cfg["dtype"] == "bf16"
    torch.manual_seed(0)
    model = _build(args.config, torch_ref=False).to(dev)
    if bf16:
        model = model.to(torch.bfloat16)
    model.train()
    opt = SGD(model.parameters(), lr=0.01, momentum=0.9, weight_decay=1e-4)
    dp = DataParallel(model, opt) if world > 1 else None
    use_graph = not args.no_graph
    trainer = Trainer(model, opt, data_parallel=dp, cuda_graph=use_graph)

    g = torch.Generator().manual_seed(1234 + rank)
    x_host = torch.randn(B, *cfg["shape"], generator=g).pin_memory()
    y_host = torch.randint(0, cfg["classes"], (B,), generator=g).pin_memory()
    x = x_host.to(dev)
    y = y_host.to(dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # Warm-up: exactly the requested number of steps (at least 3).  The CUDA-graph capture takes `graph_warmup` eager
    # steps first and happens inside these steps, so the timed region replays a captured graph from its first step.
    warm = max(args.warmup, 3)
    if use_graph:
        trainer._graph_warmup = max(1, min(trainer._graph_warmup, warm - 2))
    for _ in range(warm):
        loss = trainer.step(x, y)
    torch.cuda.synchronize()
    if use_graph and not trainer.graph_captured:
        raise SystemExit("bench.py: the step was not captured during warm-up")
    if not torch.isfinite(loss).item():
        raise SystemExit("bench.py: non-finite loss during warm-up")

    # ---------------- timed region 1: inputs resident in HBM
    prof: list = []
    if not trainer.graph_captured:
        RF.set_profile(prof)   # per-kernel CUDA-event pairs inside the timed region (eager mode)
        RF.timing_enable(True)
    clocks = ClockSampler(local)
    barrier()
    clocks.start()
    launches0 = RF.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        loss = trainer.step(x, y)
    e1.record()
    torch.cuda.synchronize()
    launches = RF.launch_count() - launches0
    barrier()
    clk = clocks.stop()
    RF.set_profile(None)
    ms = e0.elapsed_time(e1)
    t = torch.tensor([ms], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = t.item()
    value = B * world * args.steps / (ms / 1e3)
    prof_ms = ms
    if trainer.graph_captured:
        # A replayed graph has no host-visible launch boundaries, so the per-kernel event pairs are taken in a
        # second pass of the same K steps run eagerly (same kernels, same arguments, same data).
        eager = Trainer(model, opt, data_parallel=dp)
        eager.step(x, y)
        RF.set_profile(prof)
        RF.timing_enable(True)
        barrier()
        launches0 = RF.launch_count()
        # Each eager step is enqueued behind a ~30 ms spin kernel: the host gets ahead of the device, so the
        # event pairs bracket back-to-back kernels and not the Python launch path between them.
        spin = int(30e-3 * torch.cuda.get_device_properties(dev).clock_rate * 1e3) if hasattr(
            torch.cuda.get_device_properties(dev), "clock_rate") else 60_000_000
        prof_ms = 0.0
        marks = []
        for _ in range(args.steps):
            torch.cuda._sleep(spin)
            p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            p0.record()
            eager.step(x, y)
            p1.record()
            marks.append((p0, p1))
        torch.cuda.synchronize()
        RF.set_profile(None)
        prof_ms = sum(a_.elapsed_time(b_) for a_, b_ in marks)
        launches = RF.launch_count() - launches0   # the graph replays exactly these kernels, K times
    # library-side event pairs around the individual launches of the tensor-core entry points
    tc_only_ms, tc_launches = RF.timing_read("tc_kernel")
    red_ms, red_launches = RF.timing_read("splitk_reduce")
    fin_ms, fin_launches = RF.timing_read("bn_finalize")
    RF.timing_enable(False)

    # per-kernel-family device time of the profiled pass
    fam: dict[str, list] = {}
    detail: dict[tuple, list] = {}
    for kind, work, a, b, tag in prof:
        el = a.elapsed_time(b)
        d = fam.setdefault(kind, [0.0, 0.0, 0])
        d[0] += el; d[1] += (work or 0.0); d[2] += 1
        dd = detail.setdefault((kind, tag), [0.0, 0.0, 0])
        dd[0] += el; dd[1] += (work or 0.0); dd[2] += 1
    if args.detail and rank == 0:
        with open(args.detail, "w") as fh:
            fh.write("kind\tshape\tcalls_per_step\tms_per_call\twork_per_call\trate(TFLOP/s or GB/s)\n")
            for (kind, tag), (el, work, n) in sorted(detail.items(), key=lambda kv: -kv[1][0]):
                per = el / n
                w = work / n
                rate = (w / (per / 1e3) / 1e12) if kind.startswith("tc_") else (w / (per / 1e3) / 1e9)
                fh.write(f"{kind}\t{tag}\t{n / args.steps:.1f}\t{per:.4f}\t{w:.4g}\t{rate:.1f}\n")
    span_tc_ms = sum(v[0] for k, v in fam.items() if k.startswith("tc_"))
    # algorithmic FLOPs: every tensor-core launch records 2*N*K*C*R*S*P*Q (nominal channels) / 2*M*N*K
    alg_flops = sum(v[1] for k, v in fam.items() if k.startswith("tc_"))
    # the tensor-core kernel alone: the spans of the conv / GEMM entry points also cover the split-K reduction and
    # the statistics finalize kernel those entry points launch; the library-side pairs separate them
    tc_ms = tc_only_ms if tc_launches else span_tc_ms
    peaks = _peaks()
    tc_tflops = alg_flops / (tc_ms / 1e3) / 1e12 if tc_ms > 0 else 0.0
    peak_tf = peaks["bf16_tflops_sustained"]
    bn_kinds = {k: v for k, v in fam.items() if k.startswith("bn_")}
    bn_ms = sum(v[0] for v in bn_kinds.values())
    bn_bytes = sum(v[1] for v in bn_kinds.values())
    breakdown = {k: round(v[0] / args.steps, 3) for k, v in sorted(fam.items())}
    if tc_launches:
        # move the time of the non-tensor-core launches out of the tc_* spans into rows of their own
        extra = red_ms + fin_ms
        scale = (span_tc_ms - extra) / span_tc_ms if span_tc_ms > 0 else 1.0
        for k in list(breakdown):
            if k.startswith("tc_"):
                breakdown[k] = round(breakdown[k] * scale, 3)
        breakdown["splitk_reduce"] = round(red_ms / args.steps, 3)
        breakdown["bn_finalize(conv epilogue stats)"] = round(fin_ms / args.steps, 3)
    accounted = sum(v[0] for v in fam.values())
    unaccounted_ms = (prof_ms - accounted) / args.steps   # launch gaps, ATen copies / fills, NCCL

    # ---------------- timed region 2: end to end through the public API with HOST buffers
    # Every step: pinned host batch -> device (HostFeeder, copy stream, overlapped with the previous step's
    # kernels), one training step, and the step's loss read back to the host.
    from replicann_b200.training import HostFeeder
    feeder = HostFeeder(trainer, dev)
    feeder.submit(x_host, y_host)
    feeder.step()
    torch.cuda.synchronize()
    barrier()
    s0 = time.perf_counter()
    e0.record()
    last = 0.0
    feeder.submit(x_host, y_host)                       # batch of the first timed step
    for k in range(args.steps):
        if k + 1 < args.steps:
            feeder.submit(x_host, y_host)               # H2D of the next step's inputs, inside the timed region
        last = float(feeder.step().item())              # D2H of this step's loss
    e1.record()
    torch.cuda.synchronize()
    wall = time.perf_counter() - s0
    barrier()
    ms2 = max(e0.elapsed_time(e1), wall * 1e3)
    t = torch.tensor([ms2], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e = B * world * args.steps / (t.item() / 1e3)

    cpu_base = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        # a bounded sample of the same workload, sized for roughly 10 s of host work on a 16-core box
        cpu_base = _cpu_baseline(args.config, cfg["cpu_batch"], cfg["cpu_steps"], 2)

    if rank == 0:
        # DRAM bytes per launch / per step: from the committed ncu pass over one step (tools/step_profile.py)
        traffic = {}
        for name in ("r02_traffic.json", "r01_traffic.json"):
            try:
                traffic = json.loads((ROOT / "profiles" / name).read_text())
                break
            except (OSError, ValueError):
                continue
        is_r50 = args.config == "resnet50" and B == 256
        line = {
            "metric": cfg["metric"], "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": warm, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": cfg["dtype"], "data": "synthetic",
            "config": {"workload": cfg["workload"], "per_gpu_batch": B, "global_batch": B * world,
                       "parallelism": f"dp{world}", "optimizer": "SGD momentum 0.9 wd 1e-4 (fp32 master weights)",
                       "cuda_graph": bool(trainer.graph_captured),
                       "l2": "inputs larger than L2 (154 MB input batch, >10 GB of activations per step): no flush"
                       if args.config != "lenet5" else "launch-bound microsecond step; working set fits L2 by nature"},
            "e2e": {"value": e2e, "unit": UNIT,
                    "h2d_bytes_per_step": x_host.numel() * x_host.element_size() + y_host.numel() * 8,
                    "d2h_bytes_per_step": 4},
            "gpu_launches": int(launches),
            "clocks": clk,
            "roofline": {"bound": "tensor", "kernel": "tc_kernel (tcgen05 implicit-GEMM conv fprop/dgrad/wgrad + Linear GEMM)",
                         "achieved": tc_tflops, "peak": peak_tf, "unit": "TFLOP/s", "frac": tc_tflops / peak_tf,
                         "peak_source": f"{peaks['source']} bf16_tflops_sustained",
                         "traffic": traffic.get("tc_kernel_dram_bytes_per_launch") if is_r50 else None,
                         "traffic_source": traffic.get("source", "no ncu capture committed") if is_r50 else None,
                         "frac_of_nominal_2250": tc_tflops / 2250.0,
                         "launches_per_step": (tc_launches or sum(v[2] for k, v in fam.items() if k.startswith("tc_"))) / args.steps,
                         "share_of_step": tc_ms / prof_ms if prof_ms > 0 else None,
                         "timed_by": "library-side CUDA-event pairs around each tc_kernel launch (rnn_timing_*)"
                         if tc_launches else "event pairs around the conv / GEMM entry points",
                         "timed_in": "eager pass of the same K steps (graph replay has no launch boundaries)"
                         if trainer.graph_captured else "the timed region"},
            "roofline_hbm": {"bound": "hbm", "kernel": "BatchNorm stats/apply/backward family",
                             "achieved": (bn_bytes / (bn_ms / 1e3) / 1e9) if bn_ms > 0 else 0.0,
                             "peak": peaks["hbm_gbs"], "unit": "GB/s",
                             "frac": (bn_bytes / (bn_ms / 1e3) / 1e9 / peaks["hbm_gbs"]) if bn_ms > 0 else 0.0,
                             "algorithmic_bytes_per_step": bn_bytes / args.steps,
                             "dram_bytes_per_step": traffic.get("bn_family_dram_bytes_per_step") if is_r50 else None,
                             "share_of_step": bn_ms / prof_ms if prof_ms > 0 else None},
            "breakdown_ms_per_step": breakdown,
            "profiled_pass_ms_per_step": prof_ms / args.steps,
            "unaccounted_ms": round(unaccounted_ms, 3),
            "loss": last,
        }
        if cpu_base is not None:
            line["cpu_baseline"] = cpu_base
        _emit(line)
    if world > 1:
        # Tear-down of a process group whose collectives live inside a captured CUDA graph can block in the
        # NCCL destructor; everything that matters (the timed regions, the JSON line) is finished, so every
        # rank synchronises its device and leaves without running the interpreter's finalisers.
        torch.cuda.synchronize()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


# ------------------------------------------------------------------------------------------ conv sweep (configs[4])
def run_conv_sweep(args) -> None:
    """BASELINE.json configs[4]: ResNet-50 / VGG-16 layer shapes, fprop / dgrad / wgrad, each timed alone with CUDA
    events (inputs of every shape exceed nothing: they are re-used, so L2 is flushed between timed launches by
    writing a 256 MB buffer).  value = total algorithmic FLOPs / total kernel time."""
    import torch
    sys.path.insert(0, str(ROOT / "tools"))
    import conv_sweep as CS
    from replicann_b200 import functional as RF
    torch.cuda.set_device(0)
    peaks = _peaks()
    rows, tot_fl, tot_ms, n_launch = CS.sweep(RF, flush_l2=True, iters=max(args.steps, 3), warmup=max(args.warmup, 3))
    if args.detail:
        CS.write_tsv(rows, args.detail, peaks["bf16_tflops"])
    tf = tot_fl / (tot_ms / 1e3) / 1e12
    worst = min(rows, key=lambda r: r["tflops"])
    line = {"metric": "conv TFLOP/s over the ResNet-50 / VGG-16 layer shapes (fprop, dgrad, wgrad)", "value": tf,
            "unit": "TFLOP/s", "n_gpus": 1, "steps": max(args.steps, 3), "warmup": max(args.warmup, 3),
            "ms_per_step": tot_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic",
            "config": {"workload": "Conv microbench sweep: ResNet-50 (N=256) / VGG-16 (N=128) 3x3 and 1x1 layer shapes, "
                                   "fwd/dgrad/wgrad (configs[4])", "shapes": len(rows) // 3,
                       "l2": "flushed between timed launches (256 MB write)"},
            "e2e": None, "gpu_launches": n_launch,
            "roofline": {"bound": "tensor", "kernel": "tc_kernel", "achieved": tf, "peak": peaks["bf16_tflops"],
                         "unit": "TFLOP/s", "frac": tf / peaks["bf16_tflops"], "peak_source": f"{peaks['source']} bf16_tflops (burst: kernels timed alone)",
                         "frac_of_nominal_2250": tf / 2250.0, "traffic": None,
                         "worst_shape": {k: worst[k] for k in ("net", "shape", "pass", "us", "tflops")}},
            "clocks": None}
    _emit(line)


class _StdoutGuard:
    """Libraries (NCCL prints its version banner) write to fd 1; the contract is ONE JSON line on stdout.
    Everything except the final line is diverted to stderr."""

    def __enter__(self):
        sys.stdout.flush()
        self.real = os.dup(1)
        os.dup2(2, 1)
        return self

    def emit(self, line: str) -> None:
        sys.stdout.flush()
        os.write(self.real, (line + "\n").encode())

    def __exit__(self, *exc):
        sys.stdout.flush()
        os.dup2(self.real, 1)
        os.close(self.real)
        return False


_GUARD = None


def _emit(obj) -> None:
    line = json.dumps(obj)
    if _GUARD is not None:
        _GUARD.emit(line)
    else:
        print(line, flush=True)


def main() -> None:
    global _GUARD
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="resnet50", choices=list(CONFIGS) + ["conv_sweep"],
                    help="BASELINE.json config to run (default: resnet50 = configs[3], the one the metric is quoted on)")
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch (default: the config's, 256 for ResNet-50)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="run the timed steps eagerly instead of as a CUDA graph")
    ap.add_argument("--detail", default="", help="write a per-layer kernel timing table to this file")
    args = ap.parse_args()
    with _StdoutGuard() as g:
        _GUARD = g
        if args.config == "conv_sweep":
            if args.impl == "reference":
                _emit({"impl": "reference", "unavailable": "the conv sweep has no CPU arm; see --config resnet50"})
            else:
                run_conv_sweep(args)
        elif args.impl == "reference":
            run_reference(args)
        else:
            run_gpu(args)
        _GUARD = None


if __name__ == "__main__":
    main()
