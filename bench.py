#!/usr/bin/env python
"""bench.py -- BASELINE.json's headline metric on B200: train images/sec, ResNet-50 bf16, batch 256/GPU.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K --warmup W   # CPU reference arm (oracle port)

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

METRIC = "train images/sec, ResNet-50 bf16"
UNIT = "images/s"
WORKLOAD = "ResNet-50 bf16 data-parallel training, batch 256/GPU, fused BN+ReLU, synthetic 3x224x224 (configs[3])"


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
def _oracle_step_rate(batch: int, steps: int, warmup: int, threads: int):
    """Time the numpy oracle's ResNet-50 training step (the CPU restatement) on `batch` images."""
    import numpy as np
    from oracle.models_numpy import ResNetOracle
    rng = np.random.default_rng(1234)
    net = ResNetOracle((3, 4, 6, 3), True, 1000, seed=0)
    x = rng.standard_normal((batch, 3, 224, 224)).astype(np.float32)
    y = rng.integers(0, 1000, size=(batch,))
    for _ in range(warmup):
        net.train_step(x, y, lr=0.01)
    t0 = time.perf_counter()
    for _ in range(steps):
        net.train_step(x, y, lr=0.01)
    dt = time.perf_counter() - t0
    return batch * steps / dt, dt / max(steps, 1)


def run_reference(args) -> None:
    """--impl reference: the path's CPU implementation with all host threads, bounded sample per step.
    The reference repository has no CNN code and its arithmetic lives in torch, so the
    timed CPU implementation is the oracle port (numpy, multi-threaded BLAS) -- kind = "port"."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    # probe one step at batch 2 and size the per-step sample so the run stays within ~3 minutes
    _, probe = _oracle_step_rate(2, 1, 0, cores)
    per_img = probe / 2
    total_steps = args.steps + args.warmup
    batch = 8
    while batch > 1 and per_img * batch * total_steps > 170:
        batch //= 2
    rate, sec = _oracle_step_rate(batch, args.steps, args.warmup, cores)
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "per_step_sample": f"{batch} images (bounded CPU sample of the 256-image batch)"},
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{args.steps} ResNet-50 fp32 training steps of {batch} images, numpy oracle"},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    _emit(line)


# ------------------------------------------------------------------------------------------ GPU arm
def run_gpu(args) -> None:
    import torch
    import torch.distributed as dist

    from replicann_b200 import functional as RF
    from replicann_b200.arch.resnet import ResNet50
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

    B = args.batch
    torch.manual_seed(0)
    model = ResNet50().to(dev).to(torch.bfloat16)
    model.train()
    opt = SGD(model.parameters(), lr=0.01, momentum=0.9, weight_decay=1e-4)
    dp = DataParallel(model, opt) if world > 1 else None
    use_graph = not args.no_graph
    trainer = Trainer(model, opt, data_parallel=dp, cuda_graph=use_graph)

    g = torch.Generator().manual_seed(1234 + rank)
    x_host = torch.randn(B, 3, 224, 224, generator=g).pin_memory()
    y_host = torch.randint(0, 1000, (B,), generator=g).pin_memory()
    x = x_host.to(dev)
    y = y_host.to(dev)


    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    warm = max(args.warmup, 3)
    i = 0
    while i < warm or (use_graph and not trainer.graph_captured) or (use_graph and i < warm + 2):
        loss = trainer.step(x, y)   # eager steps first; with cuda_graph the capture happens here, untimed
        i += 1
    warm = i
    torch.cuda.synchronize()
    if not torch.isfinite(loss).item():
        raise SystemExit("bench.py: non-finite loss during warm-up")

    # ---------------- timed region 1: inputs resident in HBM
    prof: list = []
    if not trainer.graph_captured:
        RF.set_profile(prof)   # per-kernel CUDA-event pairs inside the timed region (eager mode)
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

    # per-kernel-family device time inside the timed region
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
    tc_ms = sum(v[0] for k, v in fam.items() if k.startswith("tc_"))
    tc_calls = sum(v[2] for k, v in fam.items() if k.startswith("tc_"))
    # algorithmic FLOPs: every tensor-core launch records 2*N*K*C*R*S*P*Q (nominal channels) / 2*M*N*K
    alg_flops = sum(v[1] for k, v in fam.items() if k.startswith("tc_"))
    peaks = _peaks()
    tc_tflops = alg_flops / (tc_ms / 1e3) / 1e12 if tc_ms > 0 else 0.0
    peak_tf = peaks["bf16_tflops_sustained"]
    hbm_kinds = {k: v for k, v in fam.items() if not k.startswith("tc_")}
    hbm_ms = sum(v[0] for v in hbm_kinds.values())
    hbm_bytes = sum(v[1] for v in hbm_kinds.values())

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
        cores = os.cpu_count() or 1
        rate, sec = _oracle_step_rate(8, 1, 0, cores)
        cpu_base = {"value": rate, "unit": UNIT, "cores": cores, "kind": "port",
                    "sample": "1 ResNet-50 fp32 training step of 8 images (numpy oracle, multi-threaded BLAS)"}

    if rank == 0:
        # DRAM bytes per tc_kernel launch: from the committed ncu pass over one step (tools/step_profile.py)
        try:
            traffic = json.loads((ROOT / "profiles" / "r01_traffic.json").read_text())
        except (OSError, ValueError):
            traffic = {}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warm,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": WORKLOAD, "per_gpu_batch": B, "global_batch": B * world, "parallelism": f"dp{world}",
                       "optimizer": "SGD momentum 0.9 wd 1e-4 (fp32 master weights)",
                       "cuda_graph": bool(trainer.graph_captured),
                       "l2": "inputs larger than L2: 154 MB input batch, >10 GB of activations per step"},
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": x_host.numel() * 4 + y_host.numel() * 8,
                    "d2h_bytes_per_step": 4},
            "gpu_launches": int(launches),
            "clocks": clk,
            "roofline": {"bound": "tensor", "kernel": "tc_kernel (tcgen05 implicit-GEMM conv fprop/dgrad/wgrad + Linear GEMM)",
                         "achieved": tc_tflops, "peak": peak_tf, "unit": "TFLOP/s", "frac": tc_tflops / peak_tf,
                         "peak_source": f"{peaks['source']} bf16_tflops_sustained",
                         "traffic": traffic.get("tc_kernel_dram_bytes_per_launch"),
                         "traffic_source": traffic.get("source", "no ncu capture committed"),
                         "frac_of_nominal_2250": tc_tflops / 2250.0,
                         "launches_per_step": tc_calls / args.steps, "share_of_step": tc_ms / prof_ms,
                         "timed_in": "eager pass of the same K steps (graph replay has no launch boundaries)"
                         if trainer.graph_captured else "the timed region"},
            "roofline_hbm": {"bound": "hbm", "kernel": "BatchNorm stats/apply/backward family",
                             "achieved": (hbm_bytes / (hbm_ms / 1e3) / 1e9) if hbm_ms > 0 else 0.0,
                             "peak": peaks["hbm_gbs"], "unit": "GB/s",
                             "frac": (hbm_bytes / (hbm_ms / 1e3) / 1e9 / peaks["hbm_gbs"]) if hbm_ms > 0 else 0.0,
                             "share_of_step": hbm_ms / prof_ms},
            "breakdown_ms_per_step": {k: round(v[0] / args.steps, 3) for k, v in sorted(fam.items())},
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
    ap.add_argument("--batch", type=int, default=256, help="per-GPU batch (256 = BASELINE.json config)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="run the timed steps eagerly instead of as a CUDA graph")
    ap.add_argument("--detail", default="", help="write a per-layer kernel timing table to this file")
    args = ap.parse_args()
    with _StdoutGuard() as g:
        _GUARD = g
        if args.impl == "reference":
            run_reference(args)
        else:
            run_gpu(args)
        _GUARD = None


if __name__ == "__main__":
    main()
