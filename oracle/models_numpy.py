"""ORACLE -- TEST INFRASTRUCTURE ONLY (see oracle/cnn_numpy.py for the rules and the pinning statement).

Whole-network CPU restatements built from the numpy operators: LeNet-5 and the ResNet family (basic and
bottleneck blocks, stride on the 3x3 convolution), forward + backward + SGD, with parameters held in a
dict keyed exactly like the state_dict of replicann_b200.arch.* so weights move across by name.
Used (a) as the whole-step parity oracle at small sizes and (b) as the timed CPU baseline of bench.py.
"""
from __future__ import annotations

import numpy as np

from . import cnn_numpy as O


def _kaiming_conv(rng, k, c, r, s, dtype):
    std = np.sqrt(2.0 / (k * r * s))  # fan_out, relu
    return (rng.standard_normal((k, c, r, s)) * std).astype(dtype)


class ResNetOracle:
    """params: '<prefix>.weight' etc.  Blocks: dicts of layer names in execution order."""

    def __init__(self, layers=(3, 4, 6, 3), bottleneck=True, n_classes=1000, base_width=64, seed=0,
                 dtype=np.float32):
        rng = np.random.default_rng(seed)
        self.dtype = dtype
        self.p: dict[str, np.ndarray] = {}
        self.blocks = []
        exp = 4 if bottleneck else 1
        self._conv("_conv1", rng, base_width, 3, 7)
        self._bn("_bn1", base_width)
        ch = base_width
        for i, n in enumerate(layers):
            w = base_width * 2 ** i
            for j in range(n):
                stride = 2 if (i > 0 and j == 0) else 1
                pre = f"_stages.{i}.{j}"
                cout = w * exp
                if bottleneck:
                    self._conv(pre + "._conv1", rng, w, ch, 1); self._bn(pre + "._bn1", w)
                    self._conv(pre + "._conv2", rng, w, w, 3); self._bn(pre + "._bn2", w)
                    self._conv(pre + "._conv3", rng, cout, w, 1); self._bn(pre + "._bn3", cout)
                else:
                    self._conv(pre + "._conv1", rng, w, ch, 3); self._bn(pre + "._bn1", w)
                    self._conv(pre + "._conv2", rng, w, w, 3); self._bn(pre + "._bn2", w)
                down = stride != 1 or ch != cout
                if down:
                    self._conv(pre + "._down_conv", rng, cout, ch, 1); self._bn(pre + "._down_bn", cout)
                self.blocks.append((pre, bottleneck, stride, down))
                ch = cout
        bound = 1 / np.sqrt(ch)
        self.p["_fc.weight"] = rng.uniform(-bound, bound, (n_classes, ch)).astype(dtype)
        self.p["_fc.bias"] = rng.uniform(-bound, bound, (n_classes,)).astype(dtype)
        self.buf = {}

    def _conv(self, name, rng, k, c, r):
        self.p[name + ".weight"] = _kaiming_conv(rng, k, c, r, r, self.dtype)

    def _bn(self, name, c):
        self.p[name + ".weight"] = np.ones(c, self.dtype)
        self.p[name + ".bias"] = np.zeros(c, self.dtype)
        self.p[name + ".running_mean"] = np.zeros(c, self.dtype)
        self.p[name + ".running_var"] = np.ones(c, self.dtype)

    def trainable(self):
        return [k for k in self.p if not k.endswith(("running_mean", "running_var"))]

    # ---- forward pieces, each pushing what backward needs onto self.tape
    def _conv_f(self, name, x, stride, pad):
        self.tape.append(("conv", name, x, stride, pad))
        return O.conv2d_fwd(x, self.p[name + ".weight"], None, stride, pad)

    def _bn_f(self, name, x, relu, residual=None):
        y, cache = O.batchnorm_fwd(x, self.p[name + ".weight"], self.p[name + ".bias"],
                                   self.p[name + ".running_mean"], self.p[name + ".running_var"])
        if residual is not None:
            y = y + residual
        if relu:
            y = O.relu_fwd(y)
        self.tape.append(("bn", name, cache, y if relu else None, residual is not None))
        return y

    def forward(self, x):
        self.tape = []
        x = x.astype(self.dtype, copy=False)
        y = self._bn_f("_bn1", self._conv_f("_conv1", x, 2, 3), True)
        yp, idx = O.maxpool2d_fwd(y, 3, 2, 1)
        self.tape.append(("maxpool", idx, y.shape))
        y = yp
        for pre, bott, stride, down in self.blocks:
            self.tape.append(("block_in", pre, down))
            xin = y
            if bott:
                t = self._bn_f(pre + "._bn1", self._conv_f(pre + "._conv1", xin, 1, 0), True)
                t = self._bn_f(pre + "._bn2", self._conv_f(pre + "._conv2", t, stride, 1), True)
                t = self._conv_f(pre + "._conv3", t, 1, 0)
                last = pre + "._bn3"
            else:
                t = self._bn_f(pre + "._bn1", self._conv_f(pre + "._conv1", xin, stride, 1), True)
                t = self._conv_f(pre + "._conv2", t, 1, 1)
                last = pre + "._bn2"
            self.tape.append(("branch",))
            idt = self._bn_f(pre + "._down_bn", self._conv_f(pre + "._down_conv", xin, stride, 0), False) if down else xin
            self.tape.append(("join",))
            y = self._bn_f(last, t, True, residual=idt)
        self.tape.append(("gap", y.shape))
        f = O.global_avgpool_fwd(y).reshape(y.shape[0], -1)
        self.tape.append(("fc", f))
        return O.linear_fwd(f, self.p["_fc.weight"], self.p["_fc.bias"])

    def backward(self, dlogits):
        """Replays the tape in reverse.  The block structure is re-derived from its markers."""
        g = {}
        tape = self.tape
        i = len(tape) - 1
        kind, f = tape[i]; i -= 1
        dx, g["_fc.weight"], g["_fc.bias"] = O.linear_bwd(f, self.p["_fc.weight"], dlogits)
        kind, shp = tape[i]; i -= 1
        d = O.global_avgpool_bwd(dx.reshape(shp[0], shp[1], 1, 1), shp)

        def bn_b(entry, d):
            _, name, cache, yrelu, has_res = entry
            if yrelu is not None:
                d = O.relu_bwd(d, yrelu)
            dxx, g[name + ".weight"], g[name + ".bias"] = O.batchnorm_bwd(d, cache)
            return dxx, (d if has_res else None)

        def conv_b(entry, d, need_dx=True):
            _, name, x, stride, pad = entry
            dxx, g[name + ".weight"], _ = O.conv2d_bwd(x, self.p[name + ".weight"], d, stride, pad, need_dx)
            return dxx

        while i >= 0:
            e = tape[i]
            if e[0] == "bn" and e[4]:  # block output: bn(last) + residual
                d_main, d_res = bn_b(e, d); i -= 1
                assert tape[i][0] == "join"; i -= 1
                # identity branch entries until "branch"
                d_idt = d_res
                while tape[i][0] != "branch":
                    ent = tape[i]
                    d_idt = bn_b(ent, d_idt)[0] if ent[0] == "bn" else conv_b(ent, d_idt)
                    i -= 1
                i -= 1  # skip "branch"
                d_t = d_main
                while tape[i][0] != "block_in":
                    ent = tape[i]
                    d_t = bn_b(ent, d_t)[0] if ent[0] == "bn" else conv_b(ent, d_t)
                    i -= 1
                i -= 1  # skip "block_in"
                d = d_t + d_idt
            elif e[0] == "maxpool":
                d = O.maxpool2d_bwd(d, e[1], e[2]); i -= 1
            elif e[0] == "bn":
                d = bn_b(e, d)[0]; i -= 1
            elif e[0] == "conv":
                d = conv_b(e, d, need_dx=False); i -= 1
            else:
                raise RuntimeError(f"unexpected tape entry {e[0]}")
        return g

    def train_step(self, x, target, lr=0.1, momentum=0.9, weight_decay=0.0):
        logits = self.forward(x)
        loss, dl = O.softmax_ce(logits, target)
        grads = self.backward(dl)
        first = not self.buf
        for k in self.trainable():
            if k not in self.buf:
                self.buf[k] = np.zeros_like(self.p[k])
            O.sgd_step(self.p[k], grads[k], self.buf[k], lr, momentum, weight_decay, first=first)
        return float(loss), logits, grads


class LeNet5Oracle:
    def __init__(self, n_classes=10, seed=0, dtype=np.float32):
        rng = np.random.default_rng(seed)
        self.dtype = dtype

        def u(shape, fan_in):
            b = 1 / np.sqrt(fan_in)
            return rng.uniform(-b, b, shape).astype(dtype)
        self.p = {
            "_conv1.weight": u((6, 1, 5, 5), 25), "_conv1.bias": u((6,), 25),
            "_conv2.weight": u((16, 6, 5, 5), 150), "_conv2.bias": u((16,), 150),
            "_fc1.weight": u((120, 400), 400), "_fc1.bias": u((120,), 400),
            "_fc2.weight": u((84, 120), 120), "_fc2.bias": u((84,), 120),
            "_fc3.weight": u((n_classes, 84), 84), "_fc3.bias": u((n_classes,), 84),
        }
        self.buf = {}

    def forward(self, x):
        p = self.p
        c = {}
        c["x"] = x.astype(self.dtype, copy=False)
        c["a1"] = O.tanh_fwd(O.conv2d_fwd(c["x"], p["_conv1.weight"], p["_conv1.bias"], 1, 2))
        c["p1"] = O.avgpool2d_fwd(c["a1"], 2)
        c["a2"] = O.tanh_fwd(O.conv2d_fwd(c["p1"], p["_conv2.weight"], p["_conv2.bias"], 1, 0))
        c["p2"] = O.avgpool2d_fwd(c["a2"], 2)
        c["f"] = c["p2"].reshape(x.shape[0], -1)
        c["h1"] = O.tanh_fwd(O.linear_fwd(c["f"], p["_fc1.weight"], p["_fc1.bias"]))
        c["h2"] = O.tanh_fwd(O.linear_fwd(c["h1"], p["_fc2.weight"], p["_fc2.bias"]))
        self.c = c
        return O.linear_fwd(c["h2"], p["_fc3.weight"], p["_fc3.bias"])

    def backward(self, dl):
        p, c, g = self.p, self.c, {}
        d, g["_fc3.weight"], g["_fc3.bias"] = O.linear_bwd(c["h2"], p["_fc3.weight"], dl)
        d = O.tanh_bwd(d, c["h2"])
        d, g["_fc2.weight"], g["_fc2.bias"] = O.linear_bwd(c["h1"], p["_fc2.weight"], d)
        d = O.tanh_bwd(d, c["h1"])
        d, g["_fc1.weight"], g["_fc1.bias"] = O.linear_bwd(c["f"], p["_fc1.weight"], d)
        d = O.avgpool2d_bwd(d.reshape(c["p2"].shape), c["a2"].shape, 2)
        d = O.tanh_bwd(d, c["a2"])
        d, g["_conv2.weight"], g["_conv2.bias"] = O.conv2d_bwd(c["p1"], p["_conv2.weight"], d, 1, 0)
        d = O.avgpool2d_bwd(d, c["a1"].shape, 2)
        d = O.tanh_bwd(d, c["a1"])
        _, g["_conv1.weight"], g["_conv1.bias"] = O.conv2d_bwd(c["x"], p["_conv1.weight"], d, 1, 2, need_dx=False)
        return g

    def train_step(self, x, target, lr=0.05, momentum=0.9):
        logits = self.forward(x)
        loss, dl = O.softmax_ce(logits, target)
        grads = self.backward(dl)
        first = not self.buf
        for k in self.p:
            if k not in self.buf:
                self.buf[k] = np.zeros_like(self.p[k])
            O.sgd_step(self.p[k], grads[k], self.buf[k], lr, momentum, first=first)
        return float(loss), logits, grads
