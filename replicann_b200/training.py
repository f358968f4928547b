"""The training-step API: forward, loss, backward, (gradient all-reduce), fused optimizer update.

The reference has no training entry point; this is the B200 form of
`loss = F.cross_entropy(model(x), y); loss.backward(); opt.step()`.
"""
from __future__ import annotations

import torch

from .nn.layers import cross_entropy


class Trainer:
    def __init__(self, model: torch.nn.Module, optimizer, *, loss_fn=cross_entropy, data_parallel=None) -> None:
        self._model, self._opt, self._loss_fn, self._dp = model, optimizer, loss_fn, data_parallel

    @property
    def model(self) -> torch.nn.Module:
        return self._model

    @property
    def optimizer(self):
        return self._opt

    def step(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        """One optimisation step on a batch already resident on the device; returns the (device) loss."""
        self._opt.zero_grad()
        if self._dp is not None:
            self._dp.begin_step()
        loss = self._loss_fn(self._model(x), y)
        loss.backward()
        if self._dp is not None:
            self._dp.finish()
        self._opt.step()
        return loss.detach()

    def step_from_host(self, x_host: torch.Tensor, y_host: torch.Tensor) -> float:
        """End-to-end step: pinned host batch -> device, train step, loss back to the host."""
        dev = next(self._model.parameters()).device
        x = x_host.to(dev, non_blocking=True)
        y = y_host.to(dev, non_blocking=True)
        return float(self.step(x, y).item())
