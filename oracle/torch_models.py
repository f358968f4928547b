"""ORACLE -- TEST INFRASTRUCTURE ONLY (rules in oracle/cnn_numpy.py): nothing under replicann_b200/ imports this.

Independent torch.nn (CPU) statements of the CNNs BASELINE.json names: standard textbook wiring written directly
against torch.nn / torch.nn.functional -- the dependency the reference's own arithmetic lives in
(pyproject.toml:11) -- with the same attribute names as replicann_b200.arch.* so that state_dicts interchange.
Used (a) as the reference side of the model-level and block-level parity tests (tests/ref_models.py re-exports
it) and (b) by bench.py's `cpu_baseline` / `--impl reference` legs as the Track-N CPU training step timed on the
host cores."""
import torch
import torch.nn as nn
import torch.nn.functional as F


class _RoundBF16(torch.autograd.Function):
    """Round to bf16 and back, in the forward AND in the backward direction."""

    @staticmethod
    def forward(ctx, x):
        return x.to(torch.bfloat16).to(torch.float32)

    @staticmethod
    def backward(ctx, g):
        return g.to(torch.bfloat16).to(torch.float32)


def ident(x):
    return x


def round_bf16(x):
    return _RoundBF16.apply(x)


# Module-level switch: `Q` is applied wherever the B200 kernels store an activation (and therefore its
# gradient) in bf16 -- after each convolution, each fused BN(+add)(+ReLU), each pool and each Linear.
# Q = ident gives the plain fp32 reference; Q = round_bf16 gives the "bf16-storage" reference that
# separates implementation bugs from the unavoidable bf16 storage noise of a deep network.
Q = ident


def set_rounding(on: bool):
    global Q
    Q = round_bf16 if on else ident


class RefBottleneck(nn.Module):
    def __init__(self, cin, width, stride=1):
        super().__init__()
        cout = width * 4
        self._conv1 = nn.Conv2d(cin, width, 1, bias=False)
        self._bn1 = nn.BatchNorm2d(width)
        self._conv2 = nn.Conv2d(width, width, 3, stride, 1, bias=False)
        self._bn2 = nn.BatchNorm2d(width)
        self._conv3 = nn.Conv2d(width, cout, 1, bias=False)
        self._bn3 = nn.BatchNorm2d(cout)
        if stride != 1 or cin != cout:
            self._down_conv = nn.Conv2d(cin, cout, 1, stride, bias=False)
            self._down_bn = nn.BatchNorm2d(cout)
        else:
            self._down_conv = None

    def forward(self, x):
        idt = x if self._down_conv is None else Q(self._down_bn(Q(self._down_conv(x))))
        y = Q(F.relu(self._bn1(Q(self._conv1(x)))))
        y = Q(F.relu(self._bn2(Q(self._conv2(y)))))
        y = self._bn3(Q(self._conv3(y)))
        return Q(F.relu(y + idt))


class RefBasicBlock(nn.Module):
    def __init__(self, cin, width, stride=1):
        super().__init__()
        self._conv1 = nn.Conv2d(cin, width, 3, stride, 1, bias=False)
        self._bn1 = nn.BatchNorm2d(width)
        self._conv2 = nn.Conv2d(width, width, 3, 1, 1, bias=False)
        self._bn2 = nn.BatchNorm2d(width)
        if stride != 1 or cin != width:
            self._down_conv = nn.Conv2d(cin, width, 1, stride, bias=False)
            self._down_bn = nn.BatchNorm2d(width)
        else:
            self._down_conv = None

    def forward(self, x):
        idt = x if self._down_conv is None else Q(self._down_bn(Q(self._down_conv(x))))
        y = Q(F.relu(self._bn1(Q(self._conv1(x)))))
        y = self._bn2(Q(self._conv2(y)))
        return Q(F.relu(y + idt))


class RefResNet(nn.Module):
    def __init__(self, layers=(3, 4, 6, 3), bottleneck=True, n_classes=1000, base_width=64):
        super().__init__()
        blk = RefBottleneck if bottleneck else RefBasicBlock
        exp = 4 if bottleneck else 1
        self._conv1 = nn.Conv2d(3, base_width, 7, 2, 3, bias=False)
        self._bn1 = nn.BatchNorm2d(base_width)
        stages, ch = [], base_width
        for i, n in enumerate(layers):
            w = base_width * 2 ** i
            bs = []
            for j in range(n):
                bs.append(blk(ch, w, 2 if (i > 0 and j == 0) else 1))
                ch = w * exp
            stages.append(nn.Sequential(*bs))
        self._stages = nn.ModuleList(stages)
        self._fc = nn.Linear(ch, n_classes)

    def forward(self, x):
        x = Q(F.relu(self._bn1(Q(self._conv1(Q(x))))))
        x = Q(F.max_pool2d(x, 3, 2, 1))   # forward no-op (values are bf16 already); rounds the gradient as the GPU does
        for s in self._stages:
            x = s(x)
        x = Q(F.adaptive_avg_pool2d(x, 1).flatten(1))
        return Q(self._fc(x))


class RefLeNet5(nn.Module):
    def __init__(self, n_classes=10):
        super().__init__()
        self._conv1 = nn.Conv2d(1, 6, 5, 1, 2)
        self._conv2 = nn.Conv2d(6, 16, 5)
        self._fc1 = nn.Linear(400, 120)
        self._fc2 = nn.Linear(120, 84)
        self._fc3 = nn.Linear(84, n_classes)

    def forward(self, x):
        x = F.avg_pool2d(torch.tanh(self._conv1(x)), 2)
        x = F.avg_pool2d(torch.tanh(self._conv2(x)), 2)
        x = x.flatten(1)
        x = torch.tanh(self._fc1(x))
        x = torch.tanh(self._fc2(x))
        return self._fc3(x)


class RefAlexNet(nn.Module):
    def __init__(self, n_classes=1000):
        super().__init__()
        self._conv1 = nn.Conv2d(3, 64, 11, 4, 2)
        self._conv2 = nn.Conv2d(64, 192, 5, 1, 2)
        self._conv3 = nn.Conv2d(192, 384, 3, 1, 1)
        self._conv4 = nn.Conv2d(384, 256, 3, 1, 1)
        self._conv5 = nn.Conv2d(256, 256, 3, 1, 1)
        self._fc1 = nn.Linear(9216, 4096)
        self._fc2 = nn.Linear(4096, 4096)
        self._fc3 = nn.Linear(4096, n_classes)

    def forward(self, x):  # dropout disabled for parity (p = 0)
        x = F.max_pool2d(Q(F.relu(self._conv1(Q(x)))), 3, 2)
        x = F.max_pool2d(Q(F.relu(self._conv2(x))), 3, 2)
        x = Q(F.relu(self._conv5(Q(F.relu(self._conv4(Q(F.relu(self._conv3(x)))))))))
        x = Q(F.adaptive_avg_pool2d(F.max_pool2d(x, 3, 2), 6)).flatten(1)
        x = Q(F.relu(self._fc1(x)))
        x = Q(F.relu(self._fc2(x)))
        return Q(self._fc3(x))


class RefVGG(nn.Module):
    def __init__(self, cfg, n_classes=1000, hidden=4096):
        super().__init__()
        feats, ch = [], 3
        for v in cfg:
            if v == "M":
                feats.append(nn.MaxPool2d(2, 2))
            else:
                feats.append(nn.Conv2d(ch, v, 3, 1, 1))
                ch = v
        self._features = nn.ModuleList(feats)
        self._fc1 = nn.Linear(ch * 49, hidden)
        self._fc2 = nn.Linear(hidden, hidden)
        self._fc3 = nn.Linear(hidden, n_classes)

    def forward(self, x):
        x = Q(x)
        for m in self._features:
            x = m(x)
            if isinstance(m, nn.Conv2d):
                x = Q(F.relu(x))
        x = Q(F.adaptive_avg_pool2d(x, 7)).flatten(1)
        x = Q(F.relu(self._fc1(x)))
        x = Q(F.relu(self._fc2(x)))
        return Q(self._fc3(x))


def load_into(ours: nn.Module, ref: nn.Module):
    """Copy ref's weights into `ours` (same keys), rounding through ours' dtypes; then write the rounded
    values back into ref so both sides start from bit-identical parameters."""
    sd = ref.state_dict()
    missing, unexpected = ours.load_state_dict(sd, strict=False)
    assert not [k for k in missing if "num_batches" not in k], missing
    assert not unexpected, unexpected
    back = {k: v.detach().float().cpu() for k, v in ours.state_dict().items() if k in sd}
    ref.load_state_dict(back, strict=False)
