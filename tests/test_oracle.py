"""CPU: pin the numpy oracle against the golden fixtures (outputs of torch CPU for the CNN operators, of the
reference's own classes for the transformer blocks -- see tests/golden/make_goldens.py)."""
from pathlib import Path

import numpy as np
import pytest

from oracle import cnn_numpy as O
from oracle import transformer_numpy as T
from oracle.models_numpy import LeNet5Oracle, ResNetOracle

G = Path(__file__).resolve().parent / "golden"


def rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return np.abs(a - b).max() / (np.abs(b).max() + 1e-12)


@pytest.fixture(scope="module")
def ops():
    return np.load(G / "cnn_ops.npz")


@pytest.mark.parametrize("tag", ["c3s1", "c3s2", "c1s2", "c7s2", "c5p2"])
def test_conv2d(ops, tag):
    g = lambda k: ops[f"conv_{tag}_{k}"]
    stride, pad = (int(v) for v in g("sp"))
    assert rel(O.conv2d_fwd(g("x"), g("w"), g("b"), stride, pad), g("y")) < 1e-5
    dx, dw, db = O.conv2d_bwd(g("x"), g("w"), g("dy"), stride, pad)
    assert rel(dx, g("dx")) < 1e-5 and rel(dw, g("dw")) < 1e-5 and rel(db, g("db")) < 1e-5


def test_batchnorm_relu_residual(ops):
    rm, rv = np.zeros(8, np.float32), np.ones(8, np.float32)
    y, cache = O.batchnorm_fwd(ops["bn_x"], ops["bn_g"], ops["bn_b"], rm, rv)
    out = O.relu_fwd(y + ops["bn_res"])
    assert rel(out, ops["bn_y"]) < 1e-5
    assert rel(rm, ops["bn_rm"]) < 1e-5 and rel(rv, ops["bn_rv"]) < 1e-5
    d = O.relu_bwd(ops["bn_dy"], out)
    dx, dg, db = O.batchnorm_bwd(d, cache)
    assert rel(dx, ops["bn_dx"]) < 1e-4 and rel(dg, ops["bn_dg"]) < 1e-5 and rel(db, ops["bn_db"]) < 1e-5
    assert rel(d, ops["bn_dres"]) < 1e-6


@pytest.mark.parametrize("tag", ["k3s2p1", "k2s2", "k3s1p1"])
def test_maxpool_indices_bit_exact(ops, tag):
    k, s, p = (int(v) for v in ops[f"maxpool_{tag}_ksp"])
    y, idx = O.maxpool2d_fwd(ops[f"maxpool_{tag}_x"], k, s, p)
    assert np.array_equal(idx, ops[f"maxpool_{tag}_idx"])
    assert np.array_equal(np.isnan(y), np.isnan(ops[f"maxpool_{tag}_y"]))
    assert np.array_equal(np.nan_to_num(y), np.nan_to_num(ops[f"maxpool_{tag}_y"]))


def test_avgpool(ops):
    x = ops["avgpool_x"]
    assert rel(O.avgpool2d_fwd(x, 2), ops["avgpool_k2"]) < 1e-6
    assert rel(O.avgpool2d_fwd(x, 3, 2, 1), ops["avgpool_k3s2p1"]) < 1e-6
    assert rel(O.global_avgpool_fwd(x), ops["avgpool_global"]) < 1e-6


def test_linear_cross_entropy(ops):
    lg = O.linear_fwd(ops["lin_x"], ops["lin_w"], ops["lin_b"])
    assert rel(lg, ops["lin_logits"]) < 1e-6
    loss, dl = O.softmax_ce(lg, ops["lin_t"])
    assert abs(loss - ops["lin_loss"]) < 1e-6
    dx, dw, db = O.linear_bwd(ops["lin_x"], ops["lin_w"], dl)
    assert rel(dx, ops["lin_dx"]) < 1e-5 and rel(dw, ops["lin_dw"]) < 1e-5 and rel(db, ops["lin_db"]) < 1e-5


def test_optimizers(ops):
    gs = ops["opt_g"]
    for name, kw in {"sgd": dict(momentum=0.9, weight_decay=1e-4), "sgd_nesterov": dict(momentum=0.9, nesterov=True)}.items():
        w, buf = ops["opt_w0"].copy(), np.zeros(64, np.float32)
        for i in range(3):
            O.sgd_step(w, gs[i], buf, 0.1, first=(i == 0), **kw)
        assert rel(w, ops[f"opt_{name}_w"]) < 1e-6, name
    w, m, v = ops["opt_w0"].copy(), np.zeros(64, np.float32), np.zeros(64, np.float32)
    for i in range(3):
        O.adamw_step(w, gs[i], m, v, i + 1, lr=1e-2, weight_decay=1e-2)
    assert rel(w, ops["opt_adamw_w"]) < 1e-6


def test_lenet5_training_step():
    g = np.load(G / "lenet5_step.npz")
    net = LeNet5Oracle()
    for k in net.p:
        net.p[k] = g["w0:" + k].copy()
    loss, logits, grads = net.train_step(g["x"], g["y"], lr=0.05, momentum=0.9)
    assert rel(logits, g["logits"]) < 1e-5 and abs(loss - g["loss"]) < 1e-5
    for k in [k[2:] for k in g.files if k.startswith("g:")]:
        assert rel(grads[k], g["g:" + k]) < 1e-4, k
    for k in [k[3:] for k in g.files if k.startswith("w1:")]:
        assert rel(net.p[k], g["w1:" + k]) < 1e-5, k


def test_resnet_tiny_training_step():
    g = np.load(G / "resnet_tiny_step.npz")
    net = ResNetOracle((1, 1), True, 10, base_width=16)
    for k in net.p:
        net.p[k] = g["w0:" + k].copy()
    loss, logits, grads = net.train_step(g["x"], g["y"], lr=0.1, momentum=0.9)
    assert rel(logits, g["logits"]) < 1e-4 and abs(loss - g["loss"]) < 1e-4
    for k in [k[2:] for k in g.files if k.startswith("g:")]:
        assert rel(grads[k], g["g:" + k]) < 1e-3, k
    assert rel(net.p["_bn1.running_var"], g["bn1_running_var"]) < 1e-5


# ------------------------------------------------------------------------------------------ Track R
@pytest.fixture(scope="module")
def tb():
    return np.load(G / "transformer_blocks.npz")


def _sd(tb, tag):
    pre = f"{tag}:sd:"
    return {k[len(pre):]: tb[k] for k in tb.files if k.startswith(pre)}


def test_encoder_and_return_kv(tb):
    assert rel(T.encoder(_sd(tb, "enc"), tb["x"], 4), tb["enc:out0"]) < 1e-5
    y, k, v = T.encoder(_sd(tb, "enc_kv"), tb["x"], 4, return_kv=True)
    assert rel(y, tb["enc_kv:out0"]) < 1e-5 and rel(k, tb["enc_kv:out1"]) < 1e-5 and rel(v, tb["enc_kv:out2"]) < 1e-5
    # return_kv skips the output projection: the two paths differ
    assert rel(tb["enc:out0"], tb["enc_kv:out0"]) > 1e-3


def test_decoder_additive_mask_is_not_causal(tb):
    p = _sd(tb, "dec")
    assert rel(T.decoder(p, tb["x"], 4), tb["dec:out0"]) < 1e-5
    x2 = tb["x"].copy()
    x2[:, -1] += 1.0  # perturb the LAST token: position 0 changes, i.e. the mask is not causal
    assert np.abs(T.decoder(p, x2, 4)[:, 0] - tb["dec:out0"][:, 0]).max() > 1e-4


def test_cross_decoder(tb):
    assert rel(T.cross_decoder(_sd(tb, "xdec"), tb["x"], tb["enc_k"], tb["enc_v"], 4), tb["xdec:out0"]) < 1e-5


def test_attention_masks(tb):
    def run(tag, mask):
        p = {"m." + k: v for k, v in _sd(tb, tag).items()}
        return T.multihead_self_attention(p, "m", tb["x"], 4, mask)
    assert rel(run("mha_bool", tb["mask_bool"]), tb["mha_bool:out0"]) < 1e-5
    assert rel(run("mha_float", tb["mask_float"]), tb["mha_float:out0"]) < 1e-5
    with pytest.raises(TypeError):
        run("mha_float", tb["mask_float"].astype(np.float64))
    p = {"m." + k: v for k, v in _sd(tb, "mca").items()}
    assert rel(T.multihead_cross_attention(p, "m", tb["x"], tb["enc_k"], tb["enc_v"], 4), tb["mca:out0"]) < 1e-5


def test_reference_cross_attention_return_kv_defect_is_recorded(tb):
    assert "not enough values to unpack" in str(tb["mca_kv_raises"])


# ------------------------------------------------------------------------------------------ stacks
@pytest.fixture(scope="module")
def ts():
    return np.load(G / "transformer_stacks.npz")


@pytest.mark.parametrize("tag", ["lm", "lm_tied_noncausal"])
def test_language_model_logits_and_loss(ts, tag):
    p = _sd(ts, tag)
    H = int(ts[f"{tag}:cfg"][2])
    logits, loss = T.language_model(p, ts[f"{tag}:ids"], H, ts[f"{tag}:targets"])
    assert rel(logits, ts[f"{tag}:logits"]) < 1e-5
    assert abs(loss - float(ts[f"{tag}:loss"])) < 1e-5 * max(1.0, abs(float(ts[f"{tag}:loss"])))
    causal = bool(ts[f"{tag}:cfg"][5])
    assert p["_decoder._blocks.0._attn_mask"].dtype == (np.bool_ if causal else np.float32)
    # causal stack: perturbing the last token leaves every earlier position's logits unchanged
    ids2 = ts[f"{tag}:ids"].copy()
    ids2[:, -1] = (ids2[:, -1] + 1) % int(ts[f"{tag}:cfg"][0])
    moved = np.abs(T.language_model(p, ids2, H)[:, :-1] - logits[:, :-1]).max()
    assert (moved < 1e-6) if causal else (moved > 1e-5)


def test_encoder_and_cross_decoder_stacks(ts):
    y, k, v = T.encoder_stack(_sd(ts, "encstack"), ts["encstack:x"], 2, return_kv=True)
    assert rel(y, ts["encstack:y"]) < 1e-5 and rel(k, ts["encstack:k"]) < 1e-5 and rel(v, ts["encstack:v"]) < 1e-5
    out = T.cross_decoder_stack(_sd(ts, "xdecstack"), ts["xdecstack:x"], k, v, 2)
    assert rel(out, ts["xdecstack:y"]) < 1e-5
