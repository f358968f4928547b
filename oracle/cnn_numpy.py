"""ORACLE -- TEST INFRASTRUCTURE ONLY.  Nothing under replicann_b200/ may import this file; only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use it, as the checker or
the timed CPU baseline, never as the product path.

CPU restatement (numpy) of the CNN training operators BASELINE.json's north_star names.  The reference
repository contains none of them; their arithmetic lives in the third-party
dependency `torch` (pyproject.toml:11 `torch = "^2.4.0"`, locked 2.4.0 at poetry.lock:1180-1181), whose
published operator semantics are restated here:

    conv2d        torch.nn.Conv2d          cross-correlation, zero padding, NCHW / KCRS
    batchnorm     torch.nn.BatchNorm2d     batch statistics, biased variance to normalise, unbiased for
                                           running_var, momentum 0.1, eps 1e-5
    maxpool2d     torch.nn.MaxPool2d       -inf padding, first maximum in row-major window order, flat
                                           h*W+w indices, NaN propagates
    avgpool2d     torch.nn.AvgPool2d       count_include_pad=True
    linear        torch.nn.Linear          x W^T + b          (reference call sites nn/attention.py:25,85)
    softmax_ce    F.cross_entropy          mean reduction
    sgd / adamw   torch.optim.SGD / AdamW

PINNING: checked against outputs of torch 2.11.0 CPU (the installed version of that dependency) frozen in
tests/golden/cnn_ops.npz by tests/golden/make_goldens.py -- see tests/test_oracle.py.  The reference
itself holds no golden vectors for any of this ("parity unpinned" by the reference).
"""
from __future__ import annotations

import numpy as np

F32 = np.float32


# ------------------------------------------------------------------------------------------ conv2d
def _im2col(x, R, S, stride, pad):
    """x [N,C,H,W] -> cols [N, C, R, S, P, Q] (a strided view of the padded input)."""
    sh, sw = stride
    ph, pw = pad
    N, C, H, W = x.shape
    xp = np.pad(x, ((0, 0), (0, 0), (ph, ph), (pw, pw))) if (ph or pw) else x
    P = (H + 2 * ph - R) // sh + 1
    Q = (W + 2 * pw - S) // sw + 1
    s = xp.strides
    cols = np.lib.stride_tricks.as_strided(xp, (N, C, R, S, P, Q), (s[0], s[1], s[2], s[3], s[2] * sh, s[3] * sw),
                                           writeable=False)
    return cols, P, Q


def _pair(v):
    return (v, v) if isinstance(v, int) else tuple(v)


def conv2d_fwd(x, w, b=None, stride=1, pad=0):
    stride, pad = _pair(stride), _pair(pad)
    K, C, R, S = w.shape
    cols, P, Q = _im2col(x, R, S, stride, pad)
    N = x.shape[0]
    y = np.einsum("kcrs,ncrspq->nkpq", w, cols, optimize=True) if N * P * Q < 4096 else \
        (w.reshape(K, -1) @ cols.reshape(N, C * R * S, P * Q)).reshape(N, K, P, Q)
    if b is not None:
        y = y + b.reshape(1, K, 1, 1)
    return y.astype(x.dtype, copy=False)


def conv2d_bwd(x, w, dy, stride=1, pad=0, need_dx=True):
    """Returns (dx, dw, db)."""
    stride, pad = _pair(stride), _pair(pad)
    sh, sw = stride
    ph, pw = pad
    K, C, R, S = w.shape
    N, _, H, W = x.shape
    cols, P, Q = _im2col(x, R, S, stride, pad)
    dy2 = dy.reshape(N, K, P * Q)
    colm = cols.reshape(N, C * R * S, P * Q)
    dw = np.zeros((K, C * R * S), dtype=x.dtype)
    for n in range(N):
        dw += dy2[n] @ colm[n].T
    dw = dw.reshape(K, C, R, S)
    db = dy.sum(axis=(0, 2, 3))
    dx = None
    if need_dx:
        dcols = (w.reshape(K, -1).T @ dy2).reshape(N, C, R, S, P, Q)  # [N, CRS, PQ]
        dxp = np.zeros((N, C, H + 2 * ph, W + 2 * pw), dtype=x.dtype)
        for r in range(R):
            for s in range(S):
                dxp[:, :, r:r + sh * P:sh, s:s + sw * Q:sw] += dcols[:, :, r, s]
        dx = dxp[:, :, ph:ph + H, pw:pw + W]
    return dx, dw, db


# ------------------------------------------------------------------------------------------ batchnorm
def batchnorm_fwd(x, gamma, beta, running_mean, running_var, momentum=0.1, eps=1e-5, training=True):
    """x [N,C,H,W]; updates running stats in place when training.  Returns (y, cache)."""
    if training:
        m = x.shape[0] * x.shape[2] * x.shape[3]
        mean = x.mean(axis=(0, 2, 3), dtype=np.float64)
        var = x.var(axis=(0, 2, 3), dtype=np.float64)
        if running_mean is not None:
            running_mean[...] = (1 - momentum) * running_mean + momentum * mean
            running_var[...] = (1 - momentum) * running_var + momentum * var * (m / max(m - 1, 1))
    else:
        mean, var = running_mean.astype(np.float64), running_var.astype(np.float64)
    invstd = 1.0 / np.sqrt(var + eps)
    xhat = (x - mean.reshape(1, -1, 1, 1)) * invstd.reshape(1, -1, 1, 1)
    y = xhat * gamma.reshape(1, -1, 1, 1) + beta.reshape(1, -1, 1, 1)
    return y.astype(x.dtype), (xhat.astype(x.dtype), invstd.astype(x.dtype), gamma)


def batchnorm_bwd(dy, cache):
    xhat, invstd, gamma = cache
    m = dy.shape[0] * dy.shape[2] * dy.shape[3]
    dbeta = dy.sum(axis=(0, 2, 3), dtype=np.float64)
    dgamma = (dy * xhat).sum(axis=(0, 2, 3), dtype=np.float64)
    g = (gamma * invstd).reshape(1, -1, 1, 1)
    dx = g * (dy - dbeta.reshape(1, -1, 1, 1) / m - xhat * dgamma.reshape(1, -1, 1, 1) / m)
    return dx.astype(dy.dtype), dgamma.astype(dy.dtype), dbeta.astype(dy.dtype)


# ------------------------------------------------------------------------------------------ activations
def relu_fwd(x):
    return np.maximum(x, 0)


def relu_bwd(dy, y):
    return dy * (y > 0)


def tanh_fwd(x):
    return np.tanh(x)


def tanh_bwd(dy, y):
    return dy * (1 - y * y)


# ------------------------------------------------------------------------------------------ pooling
def maxpool2d_fwd(x, k, stride=None, pad=0):
    """Returns (y, idx) with idx = h*W + w of the first maximum (row-major window order), as ATen."""
    k = _pair(k)
    stride = _pair(stride) if stride is not None else k
    pad = _pair(pad)
    N, C, H, W = x.shape
    ph, pw = pad
    xp = np.full((N, C, H + 2 * ph, W + 2 * pw), -np.inf, dtype=x.dtype)
    xp[:, :, ph:ph + H, pw:pw + W] = x
    cols, P, Q = _im2col(xp, k[0], k[1], stride, (0, 0))
    flat = cols.transpose(0, 1, 4, 5, 2, 3).reshape(N, C, P, Q, k[0] * k[1])
    # NaN wins, and among equal maxima the first: argmax on a key where NaN -> +inf
    key = np.where(np.isnan(flat), np.inf, flat)
    arg = key.argmax(axis=-1)
    y = np.take_along_axis(flat, arg[..., None], axis=-1)[..., 0]
    r, s = arg // k[1], arg % k[1]
    hh = np.arange(P).reshape(1, 1, P, 1) * stride[0] - ph + r
    ww = np.arange(Q).reshape(1, 1, 1, Q) * stride[1] - pw + s
    return y, (hh * W + ww).astype(np.int64)


def maxpool2d_bwd(dy, idx, x_shape):
    N, C, H, W = x_shape
    dx = np.zeros((N, C, H * W), dtype=dy.dtype)
    n = np.arange(N).reshape(N, 1, 1, 1)
    c = np.arange(C).reshape(1, C, 1, 1)
    np.add.at(dx, (np.broadcast_to(n, idx.shape), np.broadcast_to(c, idx.shape), idx), dy)
    return dx.reshape(N, C, H, W)


def avgpool2d_fwd(x, k, stride=None, pad=0):
    k = _pair(k)
    stride = _pair(stride) if stride is not None else k
    cols, P, Q = _im2col(x, k[0], k[1], stride, _pair(pad))
    return (cols.sum(axis=(2, 3)) / (k[0] * k[1])).astype(x.dtype)


def avgpool2d_bwd(dy, x_shape, k, stride=None, pad=0):
    k = _pair(k)
    stride = _pair(stride) if stride is not None else k
    ph, pw = _pair(pad)
    N, C, H, W = x_shape
    P, Q = dy.shape[2], dy.shape[3]
    dxp = np.zeros((N, C, H + 2 * ph, W + 2 * pw), dtype=dy.dtype)
    for r in range(k[0]):
        for s in range(k[1]):
            dxp[:, :, r:r + stride[0] * P:stride[0], s:s + stride[1] * Q:stride[1]] += dy / (k[0] * k[1])
    return dxp[:, :, ph:ph + H, pw:pw + W]


def global_avgpool_fwd(x):
    return x.mean(axis=(2, 3), keepdims=True).astype(x.dtype)


def global_avgpool_bwd(dy, x_shape):
    return np.broadcast_to(dy / (x_shape[2] * x_shape[3]), x_shape).astype(dy.dtype)


# ------------------------------------------------------------------------------------------ linear / loss
def linear_fwd(x, w, b=None):
    y = x @ w.T
    return y + b if b is not None else y


def linear_bwd(x, w, dy):
    return dy @ w, dy.T @ x, dy.sum(axis=0)


def softmax(x, axis=-1):
    m = x.max(axis=axis, keepdims=True)
    m = np.where(np.isfinite(m), m, 0)
    e = np.exp(x - m)
    return e / e.sum(axis=axis, keepdims=True)


def softmax_ce(logits, target):
    """Returns (mean loss, dlogits) -- F.cross_entropy with mean reduction and its gradient."""
    z = logits.astype(np.float64)
    m = z.max(axis=1, keepdims=True)
    lse = m[:, 0] + np.log(np.exp(z - m).sum(axis=1))
    rows = np.arange(z.shape[0])
    loss = (lse - z[rows, target]).mean()
    p = np.exp(z - lse[:, None])
    p[rows, target] -= 1
    return F32(loss), (p / z.shape[0]).astype(logits.dtype)


# ------------------------------------------------------------------------------------------ optimizers
def sgd_step(w, g, buf, lr, momentum=0.0, weight_decay=0.0, nesterov=False, first=False, dampening=0.0):
    """torch.optim.SGD, in place on w and buf."""
    d = g + weight_decay * w
    if momentum != 0:
        buf[...] = d if first else momentum * buf + (1 - dampening) * d
        d = d + momentum * buf if nesterov else buf
    w -= lr * d


def adamw_step(w, g, m, v, step, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2):
    """torch.optim.AdamW (amsgrad=False), in place."""
    b1, b2 = betas
    w *= 1 - lr * weight_decay
    m[...] = b1 * m + (1 - b1) * g
    v[...] = b2 * v + (1 - b2) * g * g
    bc1, bc2 = 1 - b1 ** step, 1 - b2 ** step
    w -= (lr / bc1) * m / (np.sqrt(v) / np.sqrt(bc2) + eps)
