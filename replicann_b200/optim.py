"""Fused optimizers over flat parameter buckets (torch.optim.SGD / AdamW semantics).

All parameters of one dtype live in one flat buffer; `.data` and `.grad` of every parameter are views into
it (laid out in REVERSE registration order, so gradients complete front-to-back during backward and
data-parallel buckets are contiguous ranges).  bf16 parameters additionally have an fp32 master copy that
the update kernel reads and writes together with the bf16 working copy (18 B/param for SGD-momentum).
The reference has no optimizer or training step; semantics follow torch.optim.
"""
from __future__ import annotations

from typing import Iterable

import os

import torch

from . import functional as RF

_ALIGN = 8  # elements: keeps every parameter 16-byte aligned in bf16 (TMA base-address requirement)
# Every stand-alone parameter (and the first member of a pack) starts on a 128-byte line: a filter that starts
# mid-line makes every 128-byte row of every TMA box of it straddle two lines -- invisible to kernels that load the
# filter once (resident form), a doubling of the B-operand traffic of those that stream it every K step.  Measured
# through the one kernel form that streams narrow filters (RNN_TC_DUAL=1): 685 us with a torch-allocated filter,
# 1168 us with the same filter at a 16-byte-aligned offset of the flat buffer.
_ALIGN_BASE = int(os.environ.get("RNN_PARAM_ALIGN", "64"))


class ParamPack:
    """Parameters that one kernel wants to see as ONE matrix (the per-head q/k/v projection weights of a
    multi-head attention module, nn/attention.py:85-96 of the reference keeps them as separate nn.Linear
    modules).  A module declares the pack with `ParamPack.declare([...])`; the flat optimizer then lays the
    members out back to back, in this order, so that a single view of the flat weight (and gradient) buffer is
    the packed matrix: no concatenation in forward, no split in backward, gradients land in place."""

    def __init__(self, members):
        self.members = list(members)
        self._cache = None   # (weight ptr, grad ptr, views) of the last successful lookup

    @staticmethod
    def declare(members) -> "ParamPack":
        pk = ParamPack(members)
        for m in pk.members:
            m._rnn_pack = pk
        return pk

    def views(self):
        """(packed weight view, packed gradient view) into the optimizer's flat buffers, or None when the members
        are not (or no longer) laid out contiguously there."""
        ms = self.members
        g = getattr(ms[0], "_rnn_group", None)
        if g is None:
            return None
        c = self._cache
        if c is not None:
            g0, g1 = ms[0].grad, ms[-1].grad
            if (g0 is not None and g1 is not None and c[0] == (ms[0].data_ptr(), ms[-1].data_ptr())
                    and c[1] == (g0.data_ptr(), g1.data_ptr())):
                return c[2]
        n, off0, es = ms[0].numel(), ms[0]._rnn_offset, g.work.element_size()
        for i, m in enumerate(ms):
            if getattr(m, "_rnn_group", None) is not g or m._rnn_offset != off0 + i * n or m.numel() != n:
                return None
            if m.data_ptr() != g.work.data_ptr() + (off0 + i * n) * es:
                return None   # .data was re-homed (e.g. model.to(...) after the optimizer was built)
            if m.grad is None or m.grad.data_ptr() != g.grad.data_ptr() + (off0 + i * n) * es:
                return None
        shape = (len(ms) * ms[0].shape[0],) + tuple(ms[0].shape[1:])
        total = n * len(ms)
        out = g.work[off0:off0 + total].view(shape), g.grad[off0:off0 + total].view(shape)
        self._cache = ((ms[0].data_ptr(), ms[-1].data_ptr()), (ms[0].grad.data_ptr(), ms[-1].grad.data_ptr()), out)
        return out


class _FlatGroup:
    """One dtype's parameters flattened: work (param dtype), grad (param dtype), master/state (fp32)."""

    def __init__(self, params: list[torch.nn.Parameter]):
        params, continued = self._pack_order(params)
        self.params = params
        dev, dt = params[0].device, params[0].dtype
        self.dtype = dt
        self.offsets = []
        off = 0
        for p in params:
            if id(p) not in continued:      # pack members after the first stay back to back
                off = (off + _ALIGN_BASE - 1) // _ALIGN_BASE * _ALIGN_BASE
            self.offsets.append(off)
            off += (p.numel() + _ALIGN - 1) // _ALIGN * _ALIGN
        self.numel = off
        self.work = torch.zeros(off, dtype=dt, device=dev)
        self.grad = torch.zeros(off, dtype=dt, device=dev)
        for p, o in zip(params, self.offsets):
            n = p.numel()
            self._view(self.work, p, o).copy_(p.data)
            p.data = self._view(self.work, p, o)
            p.grad = self._view(self.grad, p, o)
            p._rnn_fresh = True
            p._rnn_group = self
            p._rnn_offset = o
        self.master = self.work if dt == torch.float32 else self.work.float()
        self.state: dict[str, torch.Tensor] = {}

    @staticmethod
    def _pack_order(params: list):
        """Layout order: the given order, except that the members of a declared ParamPack sit back to back (in
        pack order) where the first of them would have been.  A pack is honoured only when all of its members
        are in this group and need no alignment padding between them."""
        ids = {id(p) for p in params}
        out, done, continued = [], set(), set()
        for p in params:
            pk = getattr(p, "_rnn_pack", None)
            if pk is None or not all(id(m) in ids and m.numel() % _ALIGN == 0 and m.numel() == pk.members[0].numel()
                                     for m in pk.members):
                if id(p) not in done:
                    out.append(p)
                    done.add(id(p))
                continue
            if id(p) in done:
                continue
            for i, m in enumerate(pk.members):
                out.append(m)
                done.add(id(m))
                if i > 0:
                    continued.add(id(m))
        return out, continued

    @staticmethod
    def _view(flat: torch.Tensor, p: torch.Tensor, off: int) -> torch.Tensor:
        n = p.numel()
        seg = flat[off:off + n]
        if p.dim() == 4:  # conv filter: logical [K,C,R,S], memory KRSC (channels_last)
            K, Cn, R, S = p.shape
            return seg.view(K, R, S, Cn).permute(0, 3, 1, 2)
        return seg.view(p.shape)

    def zeros_state(self, name: str) -> torch.Tensor:
        if name not in self.state:
            self.state[name] = torch.zeros(self.numel, dtype=torch.float32, device=self.work.device)
        return self.state[name]


class _FlatOptimizer:
    _TORCH_STATE: tuple = ()      # names of the per-parameter state tensors, as torch.optim calls them

    def __init__(self, params: Iterable[torch.nn.Parameter]):
        plist = [p for p in params if p.requires_grad]
        self._registration_order = list(plist)
        if not plist:
            raise ValueError("optimizer got an empty parameter list")
        # Bucket construction is pure bookkeeping and works on any device (the gloo tests build buckets on
        # CPU tensors); step() itself only exists as a CUDA kernel and raises on CPU parameters.
        plist = plist[::-1]  # reverse registration order ~ order in which backward produces gradients
        by_dtype: dict[torch.dtype, list] = {}
        for p in plist:
            by_dtype.setdefault(p.dtype, []).append(p)
        self.groups = [_FlatGroup(ps) for ps in by_dtype.values()]
        self.grad_scale = 1.0  # set by DataParallel to 1/world_size (all-reduce sums)
        self._steps = 0

    @property
    def flat_grads(self) -> list[torch.Tensor]:
        return [g.grad for g in self.groups]

    def zero_grad(self, set_to_none: bool = False) -> None:
        """One memset per flat gradient buffer (51 MB for ResNet-50: ~10 us), plus the `fresh` mark that lets the
        library's own gradient kernels overwrite instead of read-modify-write.  The memset is what keeps
        gradients that arrive through stock autograd correct (AccumulateGrad does `p.grad += g` into the flat
        view; e.g. a parameter used through torch.cat or any other ATen op), and it gives a parameter that
        receives no gradient in a step a zero gradient (torch.optim semantics with set_to_none=False: momentum
        and weight decay still apply).  `set_to_none` is accepted for signature parity and ignored: the
        gradients must stay views of the flat buffer."""
        self._reattach_grads()
        for g in self.groups:
            g.grad.zero_()
            for p in g.params:
                p._rnn_fresh = True

    def _reattach_grads(self) -> None:
        """`model.zero_grad()` (set_to_none=True is torch's default) or user code may have replaced a parameter's
        .grad; the fused update reads the flat buffer, so put the views back (keeping a foreign gradient's
        values)."""
        for g in self.groups:
            base, es = g.grad.data_ptr(), g.grad.element_size()
            for p, o in zip(g.params, g.offsets):
                gr = p.grad
                if gr is not None and gr.data_ptr() == base + o * es:
                    continue
                view = g._view(g.grad, p, o)
                if gr is None:
                    view.zero_()
                else:
                    view.copy_(gr)
                p.grad = view

    def state_dict(self) -> dict:
        return {"steps": self._steps,
                "groups": [{"master": g.master.clone(), **{k: v.clone() for k, v in g.state.items()}}
                           for g in self.groups]}

    def load_state_dict(self, sd: dict) -> None:
        self._steps = sd["steps"]
        for g, s in zip(self.groups, sd["groups"]):
            g.master.copy_(s["master"])
            for k, v in s.items():
                if k != "master":
                    g.zeros_state(k).copy_(v)
            if g.master is not g.work:
                RF.cast_f32_to_bf16(g.master, out=g.work)

    # ---- checkpoint interchange with torch.optim -------------------------------------
    def _located(self):
        """(registration index, parameter, group, offset) for every parameter, in registration order."""
        where = {id(p): (g, o) for g in self.groups for p, o in zip(g.params, g.offsets)}
        return [(i, p, *where[id(p)]) for i, p in enumerate(self._registration_order)]

    def _torch_hyper(self) -> dict:
        raise NotImplementedError

    def torch_state_dict(self) -> dict:
        """The optimizer state in torch.optim's own checkpoint format ({"state": {index: {...}}, "param_groups":
        [...]}, parameters numbered in registration order): per-parameter tensors are cut out of the flat fp32
        buffers and shaped like the parameter, so `torch.optim.SGD / AdamW(model.parameters()).load_state_dict()`
        accepts the result and a `.pt` written from it resumes in plain PyTorch."""
        steps = self._steps
        sd_state = {}
        for i, p, g, o in self._located():
            ent = {}
            for name in self._TORCH_STATE:
                if name in g.state:
                    ent[name] = g._view(g.state[name], p, o).detach().clone().contiguous()
            if ent or steps:
                if "step" in self._TORCH_EXTRA:
                    ent["step"] = torch.tensor(float(steps))
                if ent:
                    sd_state[i] = ent
        group = dict(self._torch_hyper(), params=list(range(len(self._registration_order))))
        return {"state": sd_state, "param_groups": [group]}

    def load_torch_state_dict(self, sd: dict) -> None:
        """Inverse of torch_state_dict(): scatter a torch.optim checkpoint into the flat buffers (hyper-parameters
        of its first param_group are adopted too).  The fp32 master weights are refreshed from the parameters."""
        pg = sd["param_groups"][0]
        self._adopt_torch_hyper(pg)
        steps = 0
        for i, p, g, o in self._located():
            ent = sd["state"].get(pg["params"][i] if i < len(pg["params"]) else i)
            if ent is None:
                continue
            for name in self._TORCH_STATE:
                if name in ent and ent[name] is not None:
                    g._view(g.zeros_state(name), p, o).copy_(ent[name].to(torch.float32))
            if "step" in ent:
                steps = max(steps, int(float(ent["step"])))
        if "step" in self._TORCH_EXTRA:
            self._steps = steps
        elif any(sd["state"].values()):
            self._steps = max(self._steps, 1)   # momentum buffers exist: the next step is not the first
        self.sync_master_from_params()

    _TORCH_EXTRA: tuple = ()

    def _adopt_torch_hyper(self, pg: dict) -> None:
        raise NotImplementedError

    def sync_master_from_params(self) -> None:
        """Call after load_state_dict() on the model: refresh the fp32 master copies from the weights."""
        for g in self.groups:
            if g.master is not g.work:
                if g.work.is_cuda:
                    RF.cast_bf16_to_f32(g.work, out=g.master)
                else:   # bookkeeping on CPU tensors (checkpoint conversion, the gloo tests): no kernel involved
                    g.master.copy_(g.work)


class SGD(_FlatOptimizer):
    _TORCH_STATE = ("momentum_buffer",)

    def _torch_hyper(self) -> dict:
        return {"lr": self.lr, "momentum": self.momentum, "dampening": self.dampening,
                "weight_decay": self.weight_decay, "nesterov": self.nesterov, "maximize": False, "foreach": None,
                "differentiable": False, "fused": None}

    def _adopt_torch_hyper(self, pg: dict) -> None:
        self.lr, self.momentum = pg.get("lr", self.lr), pg.get("momentum", self.momentum)
        self.dampening, self.weight_decay = pg.get("dampening", self.dampening), pg.get("weight_decay", self.weight_decay)
        self.nesterov = pg.get("nesterov", self.nesterov)

    def __init__(self, params, lr: float, *, momentum: float = 0.0, dampening: float = 0.0,
                 weight_decay: float = 0.0, nesterov: bool = False) -> None:
        super().__init__(params)
        if nesterov and (momentum <= 0 or dampening != 0):
            raise ValueError("Nesterov momentum requires a momentum and zero dampening")
        self.lr, self.momentum, self.dampening = lr, momentum, dampening
        self.weight_decay, self.nesterov = weight_decay, nesterov

    @torch.no_grad()
    def step(self) -> None:
        self._reattach_grads()
        first = self._steps == 0
        for g in self.groups:
            buf = g.zeros_state("momentum_buffer") if self.momentum != 0 else None
            wcopy = None if g.master is g.work else g.work
            RF.sgd_step(g.master, buf, g.grad, wcopy, lr=self.lr, momentum=self.momentum, dampening=self.dampening,
                        weight_decay=self.weight_decay, nesterov=self.nesterov, first_step=first,
                        grad_scale=self.grad_scale)
        self._steps += 1


class AdamW(_FlatOptimizer):
    _TORCH_STATE = ("exp_avg", "exp_avg_sq")
    _TORCH_EXTRA = ("step",)

    def _torch_hyper(self) -> dict:
        return {"lr": self.lr, "betas": tuple(self.betas), "eps": self.eps, "weight_decay": self.weight_decay,
                "amsgrad": False, "maximize": False, "foreach": None, "capturable": False, "differentiable": False,
                "fused": None, "decoupled_weight_decay": True}

    def _adopt_torch_hyper(self, pg: dict) -> None:
        self.lr, self.betas = pg.get("lr", self.lr), tuple(pg.get("betas", self.betas))
        self.eps, self.weight_decay = pg.get("eps", self.eps), pg.get("weight_decay", self.weight_decay)

    def torch_state_dict(self) -> dict:
        if self._step_dev is not None:
            self._steps = int(self._step_dev.item())
        return super().torch_state_dict()

    def load_torch_state_dict(self, sd: dict) -> None:
        super().load_torch_state_dict(sd)
        if self._step_dev is not None:
            self._step_dev.fill_(self._steps)

    def __init__(self, params, lr: float = 1e-3, *, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 1e-2) -> None:
        super().__init__(params)
        self.lr, self.betas, self.eps, self.weight_decay = lr, betas, eps, weight_decay
        self._step_dev = None   # int64 step count on the device: nothing host-side changes from step to step

    def _device_steps(self) -> torch.Tensor:
        if self._step_dev is None:
            self._step_dev = torch.full((1,), self._steps, dtype=torch.int64, device=self.groups[0].master.device)
        return self._step_dev

    def state_dict(self) -> dict:
        if self._step_dev is not None:
            self._steps = int(self._step_dev.item())   # replays of a captured step do not pass through step()
        return super().state_dict()

    def load_state_dict(self, sd: dict) -> None:
        super().load_state_dict(sd)
        if self._step_dev is not None:
            self._step_dev.fill_(self._steps)

    @torch.no_grad()
    def step(self) -> None:
        """The bias corrections are computed on the device from a device-side step counter, so the same sequence
        of launches serves every step (and can be replayed from a CUDA graph)."""
        self._reattach_grads()
        cnt = self._device_steps()
        self._steps += 1
        RF.counter_add(cnt, 1)
        for g in self.groups:
            wcopy = None if g.master is g.work else g.work
            RF.adamw_step(g.master, g.zeros_state("exp_avg"), g.zeros_state("exp_avg_sq"), g.grad, wcopy, lr=self.lr,
                          betas=self.betas, eps=self.eps, weight_decay=self.weight_decay, step_dev=cnt,
                          grad_scale=self.grad_scale)
