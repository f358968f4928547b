"""Data parallelism by batch: one process per GPU, bucketed gradient all-reduce over NCCL (NVLink 5 /
NVSwitch) launched from inside backward on a side stream, so the exchange overlaps the remaining
backward kernels.  The reference has no distributed code; BatchNorm statistics
stay per-rank, as torch DDP does by default.

The flat gradient buffers of the optimizer are cut into contiguous buckets (~25 MB).  Because parameters
are laid out in reverse registration order, bucket 0 completes first during backward.  When the last
gradient kernel of a bucket has been enqueued, an event is recorded on the compute stream, the
communication stream waits on it and issues `all_reduce(sum)`; the optimizer kernel then scales by
1/world_size.  Works with any torch.distributed backend (NCCL on GPUs, gloo in the CPU tests).

A bucket is launched when every DISTINCT parameter in it has reported its LAST gradient kernel of the step.  How
many gradient kernels a parameter has per step (more than one when a module is applied twice in a forward pass)
is learned during the first step, in which every bucket is launched by `finish()`; from the second step on the
launch happens inside backward.  Dropout: every rank draws from its own Philox stream (the rank is mixed into the
seed), as independent replicas would.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from .nn import _functions as Fn


class _Bucket:
    __slots__ = ("flat", "n_params", "pending", "handle", "launched")

    def __init__(self, flat: torch.Tensor, n_params: int):
        self.flat, self.n_params = flat, n_params
        self.pending, self.handle, self.launched = n_params, None, False


def plan_buckets(sizes: list[int], itemsize: int, bucket_bytes: int, tail_bytes: int = 0) -> list[tuple[int, int]]:
    """Greedy contiguous partition of parameter sizes (elements) into buckets of ~bucket_bytes.
    Returns [(first_param, last_param_exclusive)].

    `tail_bytes` > 0: the LAST bucket (the parameters whose gradients are produced last: nothing is left to overlap
    its all-reduce with) is split so that its final part is at most ~tail_bytes; only that small part's exchange is
    exposed after backward ends."""
    out, start, acc = [], 0, 0
    for i, n in enumerate(sizes):
        acc += n * itemsize
        if acc >= bucket_bytes:
            out.append((start, i + 1))
            start, acc = i + 1, 0
    if start < len(sizes):
        out.append((start, len(sizes)))
    if tail_bytes > 0 and out:
        a, b = out[-1]
        acc, cut = 0, b
        for i in range(b - 1, a, -1):       # walk back from the end while the tail stays within tail_bytes
            if acc + sizes[i] * itemsize > tail_bytes:
                break
            acc += sizes[i] * itemsize
            cut = i
        if a < cut < b:
            out[-1:] = [(a, cut), (cut, b)]
    return out


class DataParallel:
    """Wraps (model, optimizer) for synchronous data-parallel training.

    Usage per step: `dp.begin_step()`, forward, `loss.backward()`, `dp.finish()`, `optimizer.step()`.
    """

    def __init__(self, model: torch.nn.Module, optimizer, *, process_group=None, bucket_bytes: int | None = None,
                 tail_bytes: int | None = None, broadcast_parameters: bool = True) -> None:
        import os
        if bucket_bytes is None:   # RNN_DP_BUCKET_MB / RNN_DP_TAIL_MB: tuning without touching the caller
            bucket_bytes = int(float(os.environ.get("RNN_DP_BUCKET_MB", "25")) * (1 << 20))
        if tail_bytes is None:
            tail_bytes = int(float(os.environ.get("RNN_DP_TAIL_MB", "4")) * (1 << 20))
        if not dist.is_initialized():
            raise RuntimeError("torch.distributed must be initialised before building DataParallel")
        self.model, self.optimizer, self.group = model, optimizer, process_group
        self.world_size = dist.get_world_size(process_group)
        self.rank = dist.get_rank(process_group)
        optimizer.grad_scale = 1.0 / self.world_size
        self._cuda = optimizer.groups[0].work.is_cuda
        self._comm_stream = torch.cuda.Stream() if self._cuda else None
        self._buckets: list[_Bucket] = []
        self._param_bucket: dict[int, _Bucket] = {}
        for g in optimizer.groups:
            sizes = [o2 - o1 for o1, o2 in zip(g.offsets, g.offsets[1:] + [g.numel])]
            for a, b in plan_buckets(sizes, g.grad.element_size(), bucket_bytes, tail_bytes):
                lo = g.offsets[a]
                hi = g.offsets[b] if b < len(g.offsets) else g.numel
                bk = _Bucket(g.grad[lo:hi], b - a)
                self._buckets.append(bk)
                for p in g.params[a:b]:
                    self._param_bucket[id(p)] = bk
        if broadcast_parameters and self.world_size > 1:
            for g in optimizer.groups:
                dist.broadcast(g.work, src=0, group=process_group)
                if g.master is not g.work:
                    dist.broadcast(g.master, src=0, group=process_group)
            for b in model.buffers():
                dist.broadcast(b, src=0, group=process_group)
        self._uses: dict[int, int] | None = None   # gradient notifications per parameter and step (learned in step 1)
        self._count: dict[int, int] = {}
        if self.world_size > 1:
            from .nn.layers import Dropout
            Dropout._seed_mix = (self.rank * 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF
        self._hook = self._on_grad_ready
        Fn.grad_ready_hooks.append(self._hook)

    @property
    def n_buckets(self) -> int:
        return len(self._buckets)

    def close(self) -> None:
        if self._hook in Fn.grad_ready_hooks:
            Fn.grad_ready_hooks.remove(self._hook)

    def begin_step(self) -> None:
        if self._uses is None and self._count:
            self._uses = dict(self._count)
        self._count = {}
        for b in self._buckets:
            b.pending, b.handle, b.launched = b.n_params, None, False

    def _launch(self, b: _Bucket) -> None:
        b.launched = True
        if self.world_size == 1:
            return
        if self._cuda:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream())
            self._comm_stream.wait_event(ev)
            with torch.cuda.stream(self._comm_stream):
                b.handle = dist.all_reduce(b.flat, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        else:
            b.handle = dist.all_reduce(b.flat, op=dist.ReduceOp.SUM, group=self.group, async_op=True)

    def _on_grad_ready(self, p) -> None:
        b = self._param_bucket.get(id(p))
        if b is None:
            return
        c = self._count[id(p)] = self._count.get(id(p), 0) + 1
        if self._uses is None:
            return                      # first step: only learn the counts; finish() launches every bucket
        expected = self._uses.get(id(p), 1)
        if c > expected and b.launched:
            raise RuntimeError("a parameter received more gradient kernels than in the first step after its "
                               "bucket's all-reduce was launched; rebuild DataParallel for the new graph")
        if c != expected or b.launched:
            return                      # an earlier use of a shared parameter: more accumulation is to come
        b.pending -= 1
        if b.pending == 0:
            self._launch(b)

    def finish(self) -> None:
        """Launch any bucket whose parameters did not all receive gradients, then make the compute stream
        wait for every all-reduce."""
        for b in self._buckets:
            if not b.launched:
                self._launch(b)
        for b in self._buckets:
            if b.handle is not None:
                b.handle.wait()  # NCCL: makes the current stream wait on the collective; gloo: blocks
                b.handle = None
        if self._cuda and self.world_size > 1:
            torch.cuda.current_stream().wait_stream(self._comm_stream)
