"""The training-step API: forward, loss, backward, (gradient all-reduce), fused optimizer update.

The reference has no training entry point; this is the B200 form of
`loss = F.cross_entropy(model(x), y); loss.backward(); opt.step()`.

`cuda_graph=True` captures one whole step (every kernel of forward, backward and the optimizer, ~500
launches for ResNet-50) into a CUDA graph after a few eager warm-up steps and replays it afterwards: the
launch-bound gaps between the many small kernels disappear and the host does one launch per step.
"""
from __future__ import annotations

import torch

from . import functional as RF
from .nn.layers import Dropout, cross_entropy


class Trainer:
    def __init__(self, model: torch.nn.Module, optimizer, *, loss_fn=cross_entropy, data_parallel=None,
                 cuda_graph: bool = False, graph_warmup: int = 3) -> None:
        self._model, self._opt, self._loss_fn, self._dp = model, optimizer, loss_fn, data_parallel
        self._want_graph, self._graph_warmup = cuda_graph, graph_warmup
        self._graph = None
        self._eager_steps = 0
        self._has_dropout = False
        self._gx = self._gy = self._gloss = None
        self._hyper = None          # optimizer hyper-parameters baked into the captured launches
        self._bn_counts: list = []  # (BatchNorm2d, batches counted per step): replays bypass Module.forward
        self._replays = 0
        if cuda_graph:
            if graph_warmup < 1:
                # the first SGD-momentum step initialises the momentum buffer (first_step=True); captured, every
                # replay would re-initialise it
                raise ValueError("cuda_graph=True needs graph_warmup >= 1 eager step before the capture")
            self._check_graph_safe()

    @property
    def model(self) -> torch.nn.Module:
        return self._model

    @property
    def optimizer(self):
        return self._opt

    @property
    def graph_captured(self) -> bool:
        return self._graph is not None

    def _check_graph_safe(self) -> None:
        """A captured step replays identical kernel arguments, so anything that must change from step to
        step on the host side (dropout counters, Adam's bias correction, NCCL work handles) rules it out."""
        from .optim import SGD, AdamW
        if not isinstance(self._opt, (SGD, AdamW)):
            raise ValueError("cuda_graph=True supports this package's SGD and AdamW (fused, no host-side step state)")
        # With data parallelism the bucketed all-reduces are captured too: they run on the communication
        # stream, which forks from / joins the capturing stream through events (NCCL collectives are capturable).
        # Dropout is fine: the captured step starts with rnn_rng_advance(), which moves the device-side epoch every
        # Philox-drawing kernel adds to its (frozen) offset argument, so each replay draws fresh masks.
        self._has_dropout = any(isinstance(m, Dropout) or "Attention" in type(m).__name__
                                for m in self._model.modules())

    def _eager_step(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        self._opt.zero_grad()
        if self._dp is not None:
            self._dp.begin_step()
        loss = self._loss_fn(self._model(x), y)
        loss.backward()
        if self._dp is not None:
            self._dp.finish()
        self._opt.step()
        return loss.detach()

    def _opt_hyper(self) -> tuple:
        """Host scalars that become kernel arguments of the captured optimizer launches."""
        o = self._opt
        return tuple((k, getattr(o, k)) for k in ("lr", "momentum", "dampening", "weight_decay", "nesterov", "betas",
                                                  "eps", "grad_scale") if hasattr(o, k))

    def _capture(self, x: torch.Tensor, y: torch.Tensor) -> None:
        from .nn.layers import BatchNorm2d
        bns = [m for m in self._model.modules() if isinstance(m, BatchNorm2d)]
        before = [m._batches_pending for m in bns]
        self._hyper = self._opt_hyper()
        self._gx = torch.empty_like(x)
        self._gy = torch.empty_like(y)
        self._gx.copy_(x)
        self._gy.copy_(y)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        if self._dp is not None and self._dp.world_size > 1:
            # a failed NCCL capture must not take the other ranks down with a hang: relaxed capture mode lets
            # the communicator's own bookkeeping calls through
            with torch.cuda.graph(g, capture_error_mode="thread_local"):
                if self._has_dropout:
                    RF.rng_advance()
                self._gloss = self._eager_step(self._gx, self._gy)
        else:
            with torch.cuda.graph(g):
                if self._has_dropout:
                    RF.rng_advance()
                self._gloss = self._eager_step(self._gx, self._gy)
        self._graph = g
        self._bn_counts = [(m, m._batches_pending - b) for m, b in zip(bns, before) if m._batches_pending != b]
        self._replays = 0

    def step(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        """One optimisation step on a batch already resident on the device; returns the (device) loss."""
        if not self._want_graph:
            return self._eager_step(x, y)
        if self._graph is None:
            if self._eager_steps < self._graph_warmup:
                self._eager_steps += 1
                return self._eager_step(x, y)
            self._capture(x, y)
        elif x.shape != self._gx.shape or x.dtype != self._gx.dtype:
            raise ValueError("the captured step takes batches of the shape and dtype it was captured with")
        elif self._opt_hyper() != self._hyper:
            # lr / momentum / weight decay are kernel arguments frozen into the graph: a schedule that moved
            # them gets a fresh capture (one eager-speed step) instead of being silently ignored
            self._graph = None
            self._capture(x, y)
        self._gx.copy_(x, non_blocking=True)
        self._gy.copy_(y, non_blocking=True)
        self._replay()
        return self._gloss

    def _replay(self) -> None:
        self._graph.replay()
        if self._replays:   # the capture pass itself counted this step's batches once
            for m, n in self._bn_counts:
                m._batches_pending += n
        self._replays += 1

    def step_from_host(self, x_host: torch.Tensor, y_host: torch.Tensor) -> float:
        """End-to-end step: pinned host batch -> device, train step, loss back to the host."""
        if self._graph is not None:
            if self._opt_hyper() != self._hyper:
                return float(self.step(x_host.to(self._gx.device, non_blocking=True),
                                       y_host.to(self._gy.device, non_blocking=True)).item())
            self._gx.copy_(x_host, non_blocking=True)
            self._gy.copy_(y_host, non_blocking=True)
            self._replay()
            return float(self._gloss.item())
        dev = next(self._model.parameters()).device
        x = x_host.to(dev, non_blocking=True)
        y = y_host.to(dev, non_blocking=True)
        return float(self.step(x, y).item())


class HostFeeder:
    """Double-buffered host -> device input pipeline for `Trainer` (the step immediately upstream of the hot
    path).  `submit()` starts the asynchronous copy of a pinned host batch into one of
    two device staging buffers on a dedicated copy stream; `step()` makes the compute stream wait for the
    oldest submitted batch and runs one training step on it.  Submitting batch i+1 before stepping on batch i
    overlaps its PCIe transfer with the step's kernels."""

    def __init__(self, trainer: Trainer, device=None) -> None:
        self._trainer = trainer
        self._dev = device if device is not None else next(trainer.model.parameters()).device
        self._stream = torch.cuda.Stream(device=self._dev)
        self._slots: list = [None, None]      # (x_dev, y_dev, ready_event, consumed_event)
        self._w = self._r = 0
        self._pending = 0

    def submit(self, x_host: torch.Tensor, y_host: torch.Tensor) -> None:
        if self._pending >= 2:
            raise RuntimeError("HostFeeder holds two batches already: call step() before submitting more")
        i = self._w
        slot = self._slots[i]
        if slot is None or slot[0].shape != x_host.shape or slot[0].dtype != x_host.dtype:
            slot = [torch.empty(x_host.shape, dtype=x_host.dtype, device=self._dev),
                    torch.empty(y_host.shape, dtype=y_host.dtype, device=self._dev),
                    torch.cuda.Event(), None]
            self._slots[i] = slot
        with torch.cuda.stream(self._stream):
            if slot[3] is not None:
                self._stream.wait_event(slot[3])   # the step that last read this slot has finished with it
            slot[0].copy_(x_host, non_blocking=True)
            slot[1].copy_(y_host, non_blocking=True)
            slot[2].record(self._stream)
        self._w ^= 1
        self._pending += 1

    def step(self) -> torch.Tensor:
        if self._pending == 0:
            raise RuntimeError("HostFeeder.step() without a submitted batch")
        slot = self._slots[self._r]
        cur = torch.cuda.current_stream(self._dev)
        cur.wait_event(slot[2])
        loss = self._trainer.step(slot[0], slot[1])
        ev = torch.cuda.Event()
        ev.record(cur)
        slot[3] = ev
        self._r ^= 1
        self._pending -= 1
        return loss
