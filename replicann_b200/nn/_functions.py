"""torch.autograd.Function wrappers: one per operator, forward and backward both running the library's
CUDA kernels.  The reference has no custom Functions ("backward of all of the above
-- implicit autograd"); these replace ATen's forward AND derivative kernels for the same operators.

Layout convention: 4-D activations are logical NCHW tensors with channels_last strides, so
`x.permute(0, 2, 3, 1)` is the contiguous NHWC buffer the kernels read.  Parameter gradients are written
straight into `param.grad` (which may be a view into a flat optimizer bucket) and reported to
`grad_ready_hooks` so that data-parallel buckets can start their all-reduce during backward.
"""
from __future__ import annotations

import torch
from torch.autograd import Function

from .. import functional as RF

# callables invoked as hook(param) right after a parameter's gradient kernel has been enqueued
grad_ready_hooks: list = []


def _notify(p):
    for h in grad_ready_hooks:
        h(p)


def grad_target(p: torch.Tensor):
    """Tensor to write p's gradient into and whether the write must accumulate."""
    if p.grad is None:
        p.grad = torch.empty_like(p)
        return p.grad, False
    if getattr(p, "_rnn_fresh", False):
        p._rnn_fresh = False
        return p.grad, False
    return p.grad, True


def nhwc(x: torch.Tensor) -> torch.Tensor:
    """Contiguous NHWC buffer behind a logical NCHW tensor (a view when x is channels_last)."""
    if x.dim() != 4:
        raise ValueError(f"expected a 4-D NCHW tensor, got shape {tuple(x.shape)}")
    xp = x.permute(0, 2, 3, 1)
    return xp if xp.is_contiguous() else xp.contiguous()


def nchw(y_nhwc: torch.Tensor) -> torch.Tensor:
    return y_nhwc.permute(0, 3, 1, 2)


def _bias_f32(b):
    if b is None:
        return None
    if b.dtype == torch.float32:
        return b.detach()
    n = b.numel()
    if n % 4 == 0:
        return RF.cast_bf16_to_f32(b.detach().contiguous())
    return b.detach().float()


def _write_param_grad(p, compute):
    """compute(out, accumulate) enqueues the kernel(s) that produce p's gradient.

    Leaf parameters get the result written straight into `p.grad` (None is returned to autograd);
    any other tensor (e.g. a padded copy of a parameter) gets a fresh gradient tensor back."""
    if p.is_leaf:
        tgt, acc = grad_target(p)
        compute(tgt, acc)
        _notify(p)
        return None
    out = torch.empty_like(p)
    compute(out, False)
    return out


# ------------------------------------------------------------------------------------------ Conv2d
class Conv2dFn(Function):
    """y = [relu](conv2d(x, w) + bias).

    Two ways of moving the fused ReLU's backward out of this node (both declared by the model, see
    nn.layers.fuse_relu_pool / fuse_relu_conv; without them the backward pass masks dy itself):
      premasked: the only consumer of y is a pooling layer that masks the gradient at the pooled resolution;
      emit_mask: the only consumer is a unit-stride convolution -- the forward also returns the 1-bit mask (y > 0),
                 which that convolution receives as `in_mask` and applies in the epilogue of its dgrad kernel.
    Either way this node's incoming gradient is already masked and the bias gradient is a plain column sum."""

    @staticmethod
    def forward(ctx, x, weight, bias, stride, pad, relu=False, premasked=False, in_mask=None, emit_mask=False):
        xh = nhwc(x)
        wk = nhwc(weight)  # [K,C,R,S] channels_last -> KRSC
        ctx.stride, ctx.pad = stride, pad
        ctx.relu = relu and not premasked and not emit_mask
        mask = None
        if weight.dtype == torch.float32:
            if in_mask is not None or emit_mask:
                raise RuntimeError("1-bit ReLU masks are a bf16-path feature")
            y = RF.conv2d_fprop_f32(xh, wk, bias.detach() if bias is not None else None, stride=stride, pad=pad)
            if relu:
                y = RF.relu_fwd(y)
        elif emit_mask:
            y, mask = RF.conv2d_fprop_relu_mask(xh, wk, stride=stride, pad=pad, shift=_bias_f32(bias))
        else:
            y = RF.conv2d_fprop(xh, wk, stride=stride, pad=pad, shift=_bias_f32(bias), relu=relu)
        ctx.save_for_backward(xh, weight, bias, y if ctx.relu else None, in_mask)
        if emit_mask:
            ctx.mark_non_differentiable(mask)
            return nchw(y), mask
        return nchw(y)

    @staticmethod
    def backward(ctx, dy, *_):
        xh, weight, bias, y, in_mask = ctx.saved_tensors
        wk = nhwc(weight)
        dyh = nhwc(dy)
        stride, pad = ctx.stride, ctx.pad
        f32 = weight.dtype == torch.float32
        dw = db = dx = None
        want_db = bias is not None and ctx.needs_input_grad[2]
        if ctx.relu and want_db and bias.is_leaf:
            # ReLU backward and the bias gradient in one pass over dy
            box = []
            _write_param_grad(bias, lambda out, acc: box.append(RF.relu_bwd_colsum(dyh, y, out=out, accumulate=acc)))
            dyh = box[0]
            want_db = False
        elif ctx.relu:
            dyh = RF.relu_bwd(dyh, y)
        if ctx.needs_input_grad[1]:
            def wg(out, acc):
                o = nhwc_param_view(out)
                if f32:
                    RF.conv2d_wgrad_f32(xh, dyh, tuple(wk.shape), stride=stride, pad=pad, out=o, accumulate=acc)
                else:
                    RF.conv2d_wgrad(xh, dyh, tuple(wk.shape), stride=stride, pad=pad, out=o, accumulate=acc)
            dw = _write_param_grad(weight, wg)
        if want_db:
            K = dyh.shape[-1]
            db = _write_param_grad(bias, lambda out, acc: RF.colsum(dyh.reshape(-1, K), out=out, accumulate=acc))
        if ctx.needs_input_grad[0]:
            if f32:
                dx = nchw(RF.conv2d_dgrad_f32(dyh, wk, tuple(xh.shape), stride=stride, pad=pad))
            else:
                dx = nchw(RF.conv2d_dgrad(dyh, wk, tuple(xh.shape), stride=stride, pad=pad, relu_mask=in_mask))
        return dx, dw, db, None, None, None, None, None, None


class PadFilterFn(Function):
    """[K, C, R, S] filter -> [K, Cpad, R, S] (channels_last) with zeros in the added input channels: lets a
    3-channel first layer ride the 64-channel K step of the tensor-core kernel on a channel-padded input."""

    @staticmethod
    def forward(ctx, w, cpad):
        K, Cn, R, S = w.shape
        out = torch.zeros((K, cpad, R, S), dtype=w.dtype, device=w.device).contiguous(memory_format=torch.channels_last)
        out[:, :Cn].copy_(w)
        ctx.c = Cn
        return out

    @staticmethod
    def backward(ctx, g):
        return g[:, :ctx.c].contiguous(memory_format=torch.channels_last), None


class FirstConv2dFn(Function):
    """First layer of a bf16 CNN applied directly to fp32 NCHW images (C <= 8): the layout/cast kernel, the
    packed-pixel tcgen05 convolution (see rnn_conv2d_first_fprop_bf16) and its filter gradient.  No input
    gradient is produced (images do not require grad)."""

    @staticmethod
    def forward(ctx, x, weight, bias, stride, pad, relu, norm=None, premasked=False, emit_mask=False):
        from .._lib import conv_desc
        N, Cn, H, W = x.shape
        K, _, R, S = weight.shape
        d = conv_desc(N, H, W, Cn, K, R, S, stride, pad)
        _, _, sg = RF.first_conv_geometry(d)
        mean, inv_std = norm if norm is not None else (None, None)
        xp = RF.pack_first_input(x if x.is_contiguous() else x.contiguous(), d, mean, inv_std)
        # filter [K,C,R,S] -> [K,R,SG*8 taps,8 ch] with zeros beyond S and C (tiny; data movement only)
        wp = torch.zeros((K, R, sg * 8, 8), dtype=weight.dtype, device=weight.device)
        wp[:, :, :S, :Cn].copy_(weight.detach().permute(0, 2, 3, 1))
        mask = None
        if emit_mask:
            y, mask = RF.first_conv_fprop(xp, wp.reshape(K, R, sg, 64), d, shift=_bias_f32(bias), relu_mask=True)
        else:
            y = RF.first_conv_fprop(xp, wp.reshape(K, R, sg, 64), d, shift=_bias_f32(bias), relu=relu)
        ctx.d, ctx.sg, ctx.relu = d, sg, relu and not premasked and not emit_mask   # see Conv2dFn
        ctx.save_for_backward(xp, weight, bias, y if ctx.relu else None)
        if emit_mask:
            ctx.mark_non_differentiable(mask)
            return nchw(y), mask
        return nchw(y)

    @staticmethod
    def backward(ctx, dy, *_):
        xp, weight, bias, y = ctx.saved_tensors
        d, sg = ctx.d, ctx.sg
        dyh = nhwc(dy)
        dw = db = None
        want_db = bias is not None and ctx.needs_input_grad[2]
        if ctx.relu and want_db and bias.is_leaf:
            box = []   # ReLU backward and the bias gradient in one pass over dy
            _write_param_grad(bias, lambda out, acc: box.append(RF.relu_bwd_colsum(dyh, y, out=out, accumulate=acc)))
            dyh = box[0]
            want_db = False
        elif ctx.relu:
            dyh = RF.relu_bwd(dyh, y)
        if ctx.needs_input_grad[1]:
            def wg(out, acc):
                g = RF.first_conv_wgrad(xp, dyh, d, sg)
                gv = g.reshape(d.K, d.R, sg * 8, 8)[:, :, :d.S, :d.C].permute(0, 3, 1, 2)
                if acc:
                    out.add_(gv)
                else:
                    out.copy_(gv)
            dw = _write_param_grad(weight, wg)
        if want_db:
            K = dyh.shape[-1]
            db = _write_param_grad(bias, lambda out, acc: RF.colsum(dyh.reshape(-1, K), out=out, accumulate=acc))
        return None, dw, db, None, None, None, None, None, None


def nhwc_param_view(g: torch.Tensor) -> torch.Tensor:
    """KRSC buffer behind a [K,C,R,S] gradient tensor; it must already be channels_last (no copy allowed,
    the kernel writes through it)."""
    v = g.permute(0, 2, 3, 1)
    if not v.is_contiguous():
        raise RuntimeError("conv weight gradients must be channels_last (use replicann_b200.nn.Conv2d parameters)")
    return v


# ------------------------------------------------------------------------------------------ Linear
class LinearFn(Function):
    """y = [relu](x W^T + b).  The bf16 tensor-core GEMM stores 16-byte pieces of output rows, so it wants
    out_features % 8 == 0; a layer that is not (a 10- or 100-class head) runs with its weight, bias and output
    gradient zero-padded to the next multiple of 8 -- a few KB of copies on a layer that small -- and the
    padding columns are dropped again."""

    @staticmethod
    def forward(ctx, x, weight, bias, relu):
        x2 = x.reshape(-1, x.shape[-1])
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        N = weight.shape[0]
        npad = (-N) % 8 if weight.dtype == torch.bfloat16 else 0
        w = weight.detach()
        b = bias.detach() if bias is not None else None
        if npad:
            w = torch.cat([w, w.new_zeros((npad, w.shape[1]))], dim=0)
            b = torch.cat([b, b.new_zeros(npad)]) if b is not None else None
        if weight.dtype == torch.float32:
            y = RF.gemm_f32(x2, w, bias=b, relu=relu)
        else:
            y = RF.gemm(x2, w, bias=_bias_f32(b), relu=relu)
        ctx.relu, ctx.npad = relu, npad
        ctx.xshape = x.shape
        ctx.save_for_backward(x2, weight, bias, y if relu else None, w if npad else None)
        if npad:
            y = y[:, :N].contiguous()
        return y.reshape(*x.shape[:-1], N)

    @staticmethod
    def backward(ctx, dy):
        x2, weight, bias, y, wpad = ctx.saved_tensors
        N, npad = weight.shape[0], ctx.npad
        dy2 = dy.reshape(-1, N)
        if npad:
            dy2 = torch.cat([dy2, dy2.new_zeros((dy2.shape[0], npad))], dim=1)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        f32 = weight.dtype == torch.float32
        wmat = wpad if npad else weight.detach()
        dw = db = dx = None
        want_db = bias is not None and ctx.needs_input_grad[2]
        if ctx.relu and want_db and bias.is_leaf and not npad:
            box = []
            _write_param_grad(bias, lambda out, acc: box.append(RF.relu_bwd_colsum(dy2, y, out=out, accumulate=acc)))
            dy2 = box[0]
            want_db = False
        elif ctx.relu:
            dy2 = RF.relu_bwd(dy2, y)
        if ctx.needs_input_grad[1]:
            def wg(out, acc):
                if npad:   # padded rows of the gradient are dropped
                    g = RF.gemm(dy2, x2, a_mn=True, b_mn=True)[:N]
                    (out.add_ if acc else out.copy_)(g)
                elif f32:
                    RF.gemm_f32(dy2, x2, a_mn=True, b_mn=True, out=out, accumulate=acc)
                elif not acc:
                    RF.gemm(dy2, x2, a_mn=True, b_mn=True, out=out)
                else:
                    g = RF.gemm(dy2, x2, a_mn=True, b_mn=True)
                    RF.add(out, g, out=out)
            dw = _write_param_grad(weight, wg)
        if want_db:
            def bg(out, acc):
                if npad:
                    g = RF.colsum(dy2)[:N]
                    (out.add_ if acc else out.copy_)(g)
                else:
                    RF.colsum(dy2, out=out, accumulate=acc)
            db = _write_param_grad(bias, bg)
        if ctx.needs_input_grad[0]:
            if f32:
                dx = RF.gemm_f32(dy2, wmat, b_mn=True)
            else:
                dx = RF.gemm(dy2, wmat, b_mn=True)
            dx = dx.reshape(ctx.xshape)
        return dx, dw, db, None


class PackedLinearFn(Function):
    """y = x . W^T (+ b) where W stacks the weights of several nn.Linear modules row-wise -- the per-head q / k / v
    projections of a multi-head attention module (reference nn/attention.py:85-96, 149-192 runs them one head at a
    time and concatenates) as ONE tcgen05 GEMM.

    `pack` = (W view, dW view, b view | None, db view | None) into the flat optimizer buffers when the member
    parameters are laid out back to back there (optim.ParamPack): nothing is concatenated and the weight gradient
    is written in place.  `pack` = None: W is assembled with one torch.cat of the members and the gradient is
    scattered back into each member's .grad (honouring the optimizer's overwrite / accumulate state), so either way
    no gradient reaches the members through autograd's AccumulateGrad.
    The member parameters are passed as inputs only so that the node exists when x itself needs no gradient."""

    @staticmethod
    def forward(ctx, x, relu, n_w, pack, *members):
        ws, bs = members[:n_w], members[n_w:]
        if pack is not None:
            w, gw, b, gb = pack
        else:
            w = torch.cat([m.detach() for m in ws], dim=0)
            b = torch.cat([m.detach() for m in bs], dim=0) if bs else None
            gw = gb = None
        x2 = x.reshape(-1, x.shape[-1])
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        if w.dtype == torch.float32:
            y = RF.gemm_f32(x2, w, bias=b, relu=relu)
        else:
            y = RF.gemm(x2, w, bias=_bias_f32(b), relu=relu)
        ctx.relu, ctx.xshape, ctx.n_w = relu, x.shape, n_w
        ctx.members, ctx.gw, ctx.gb = members, gw, gb
        ctx.save_for_backward(x2, w, y if relu else None)
        return y.reshape(*x.shape[:-1], w.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, w, y = ctx.saved_tensors
        ws, bs = ctx.members[:ctx.n_w], ctx.members[ctx.n_w:]
        dy2 = dy.reshape(-1, w.shape[0])
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        if ctx.relu:
            dy2 = RF.relu_bwd(dy2, y)
        f32 = w.dtype == torch.float32

        def state(ms):
            """overwrite / accumulate state of a member list: all members of a pack move together"""
            acc = [grad_target(m)[1] for m in ms]
            if any(a != acc[0] for a in acc):
                raise RuntimeError("packed parameters have diverging gradient state")
            return acc[0]

        def wgrad(out, acc):
            if f32:
                RF.gemm_f32(dy2, x2, a_mn=True, b_mn=True, out=out, accumulate=acc)
            elif not acc:
                RF.gemm(dy2, x2, a_mn=True, b_mn=True, out=out)
            else:
                RF.add(out, RF.gemm(dy2, x2, a_mn=True, b_mn=True), out=out)

        if ws[0].requires_grad:
            if ctx.gw is not None:
                wgrad(ctx.gw, state(ws))
            else:
                g = torch.empty_like(w)
                wgrad(g, False)
                n = ws[0].shape[0]
                for i, m in enumerate(ws):
                    tgt, acc = grad_target(m)
                    (tgt.add_ if acc else tgt.copy_)(g[i * n:(i + 1) * n])
            for m in ws:
                _notify(m)
        if bs and bs[0].requires_grad:
            if ctx.gb is not None:
                RF.colsum(dy2, out=ctx.gb, accumulate=state(bs))
            else:
                g = RF.colsum(dy2)
                n = bs[0].shape[0]
                for i, m in enumerate(bs):
                    tgt, acc = grad_target(m)
                    (tgt.add_ if acc else tgt.copy_)(g[i * n:(i + 1) * n])
            for m in bs:
                _notify(m)
        dx = None
        if ctx.needs_input_grad[0]:
            dx = (RF.gemm_f32(dy2, w, b_mn=True) if f32 else RF.gemm(dy2, w, b_mn=True)).reshape(ctx.xshape)
        return (dx, None, None, None) + (None,) * len(ctx.members)


# ------------------------------------------------------------------------------------------ BatchNorm2d
class BatchNormFn(Function):
    """y = relu(BN(x) + residual) with batch statistics (training) or running statistics (eval)."""

    @staticmethod
    def forward(ctx, x, gamma, beta, residual, running_mean, running_var, training, momentum, eps, relu):
        xh = nhwc(x)
        Cn = xh.shape[-1]
        x2 = xh.reshape(-1, Cn)
        r2 = nhwc(residual).reshape(-1, Cn) if residual is not None else None
        g = gamma.detach() if gamma is not None else None
        b = beta.detach() if beta is not None else None
        if training:
            mean, invstd, scale, shift = RF.bn_stats(x2, g, b, eps, momentum, running_mean, running_var)
        else:
            scale, shift = RF.bn_eval_prepare(g, b, running_mean, running_var, eps)
            mean, invstd = running_mean, torch.rsqrt(running_var + eps)
        # ReLU mask for backward: recomputed from x with the forward's scale/shift when no residual was added,
        # otherwise a 1-bit-per-element mask written by the apply kernel (16x less traffic than re-reading y)
        from_x = relu and residual is None
        want_mask = relu and not from_x
        out = RF.bn_apply(x2, scale, shift, r2, relu, want_mask=want_mask)
        y2, mask = out if want_mask else (out, None)
        ctx.relu, ctx.training, ctx.has_res = relu, training, residual is not None
        ctx.save_for_backward(x2, mask, mean, invstd, gamma, beta,
                              scale if from_x else None, shift if from_x else None)
        ctx.shape = xh.shape
        return nchw(y2.reshape(xh.shape))

    @staticmethod
    def backward(ctx, dy):
        x2, mask, mean, invstd, gamma, beta, fscale, fshift = ctx.saved_tensors
        if not ctx.training:
            raise RuntimeError("BatchNorm2d backward in eval mode is not supported by replicann_b200")
        Cn = x2.shape[1]
        dy2 = nhwc(dy).reshape(-1, Cn)
        need_g = gamma is not None and gamma.requires_grad
        if need_g:
            dg, acc = grad_target(gamma)
            db, acc_b = grad_target(beta)
            if acc != acc_b:
                raise RuntimeError("BatchNorm weight/bias gradient state out of sync")
        else:
            dg = db = None
            acc = False
        dx2, dres2 = RF.bn_bwd(dy2, x2, None, mean, invstd, gamma.detach() if gamma is not None else None,
                               want_dres=ctx.has_res, dgamma=dg, dbeta=db, accumulate=acc, fwd_scale=fscale,
                               fwd_shift=fshift, relu_mask=mask)
        if need_g:
            _notify(gamma)
            _notify(beta)
        dx = nchw(dx2.reshape(ctx.shape)) if ctx.needs_input_grad[0] else None
        dres = nchw(dres2.reshape(ctx.shape)) if ctx.has_res else None
        return dx, None, None, dres, None, None, None, None, None, None


# ------------------------------------------------------------------------------------------ LayerNorm
class LayerNormFn(Function):
    """y = LayerNorm(x + residual) * gamma + beta  (post-LN block, arch/transformer.py:111-112)."""

    @staticmethod
    def forward(ctx, x, residual, gamma, beta, eps):
        E = x.shape[-1]
        x2 = x.reshape(-1, E)
        x2 = x2 if x2.is_contiguous() else x2.contiguous()
        r2 = None
        if residual is not None:
            r2 = residual.reshape(-1, E)
            r2 = r2 if r2.is_contiguous() else r2.contiguous()
        y2, xin, mean, rstd = RF.layernorm_fwd(x2, r2, gamma.detach(), beta.detach(), eps)
        ctx.has_res = residual is not None
        ctx.shape = x.shape
        ctx.save_for_backward(xin, mean, rstd, gamma, beta)
        return y2.reshape(x.shape)

    @staticmethod
    def backward(ctx, dy):
        xin, mean, rstd, gamma, beta = ctx.saved_tensors
        E = xin.shape[1]
        dy2 = dy.reshape(-1, E)
        dy2 = dy2 if dy2.is_contiguous() else dy2.contiguous()
        dg, acc = grad_target(gamma)
        db, acc_b = grad_target(beta)
        if acc != acc_b:
            raise RuntimeError("LayerNorm weight/bias gradient state out of sync")
        dx2 = RF.layernorm_bwd(dy2, xin, gamma.detach(), mean, rstd, dg, db, acc)
        _notify(gamma)
        _notify(beta)
        dx = dx2.reshape(ctx.shape)
        return dx, (dx if ctx.has_res else None), None, None, None


# ------------------------------------------------------------------------------------------ activations
class ReLUFn(Function):
    @staticmethod
    def forward(ctx, x):
        if x.dim() == 4:
            xh = nhwc(x)
            y = RF.relu_fwd(xh)
            ctx.save_for_backward(y)
            ctx.four = True
            return nchw(y)
        xc = x if x.is_contiguous() else x.contiguous()
        y = RF.relu_fwd(xc)
        ctx.save_for_backward(y)
        ctx.four = False
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        if ctx.four:
            return nchw(RF.relu_bwd(nhwc(dy), y))
        return RF.relu_bwd(dy if dy.is_contiguous() else dy.contiguous(), y)


class TanhFn(Function):
    @staticmethod
    def forward(ctx, x):
        four = x.dim() == 4
        xc = nhwc(x) if four else (x if x.is_contiguous() else x.contiguous())
        y = RF.tanh_fwd(xc)
        ctx.save_for_backward(y)
        ctx.four = four
        return nchw(y) if four else y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        if ctx.four:
            return nchw(RF.tanh_bwd(nhwc(dy), y))
        return RF.tanh_bwd(dy if dy.is_contiguous() else dy.contiguous(), y)


class DropoutFn(Function):
    @staticmethod
    def forward(ctx, x, p, seed, offset):
        xc = x if x.is_contiguous() else x.contiguous()
        y, mask = RF.dropout_fwd(xc, p, seed, offset)
        ctx.p = p
        ctx.save_for_backward(mask)
        return y

    @staticmethod
    def backward(ctx, dy):
        (mask,) = ctx.saved_tensors
        return RF.dropout_bwd(dy if dy.is_contiguous() else dy.contiguous(), mask, ctx.p), None, None, None


# ------------------------------------------------------------------------------------------ pooling
class MaxPool2dFn(Function):
    """Returns (y, indices).  `indices` (ATen's flat argmax) is only produced on request; the backward pass
    uses the 1-byte in-window argmax."""

    @staticmethod
    def forward(ctx, x, k, s, p, want_indices, relu_input=False):
        # relu_input: x is the output of a ReLU whose backward is left to this layer.  A window's gradient goes to its
        # argmax element, whose value IS the pooled value, so "x > 0 there" == "y > 0": the ReLU mask is applied to dy
        # at the pooled resolution (1/4 .. 1/2.25 of the elements) instead of in a pass over the full-size tensor.
        xh = nhwc(x)
        y, idx, widx = RF.maxpool2d_fwd(xh, k, s, p, return_indices=want_indices, return_window_idx=True)
        ctx.cfg = (tuple(xh.shape), k, s, p)
        ctx.save_for_backward(widx, y if relu_input else None)
        if idx is None:
            idx = torch.empty(0, dtype=torch.int32, device=x.device)
            ctx.mark_non_differentiable(idx)
            return nchw(y), idx
        ctx.mark_non_differentiable(idx)
        return nchw(y), nchw(idx)

    @staticmethod
    def backward(ctx, dy, _didx):
        widx, y = ctx.saved_tensors
        shape, k, s, p = ctx.cfg
        return (nchw(RF.maxpool2d_bwd(nhwc(dy), None, shape, k, s, p, window_idx=widx, relu_y=y)),
                None, None, None, None, None)


class BnReluMaxPoolFn(Function):
    """max_pool(relu(BatchNorm(x))) -- the stem of the ResNets -- with the BatchNorm-apply + ReLU folded into the
    pooling kernel: the normalised 112 x 112 activation (411 MB at batch 256) is neither written nor re-read.
    Values, argmax ties and all gradients are those of BatchNorm2d(relu=True) followed by MaxPool2d: the pooled
    values are rounded to bf16 before they are compared, and backward is the same two kernels (pool backward by
    window index, BatchNorm backward with the ReLU mask recomputed from x)."""

    @staticmethod
    def forward(ctx, x, gamma, beta, running_mean, running_var, training, momentum, eps, k, s, p):
        xh = nhwc(x)
        Cn = xh.shape[-1]
        x2 = xh.reshape(-1, Cn)
        g, b = gamma.detach(), beta.detach()
        if training:
            mean, invstd, scale, shift = RF.bn_stats(x2, g, b, eps, momentum, running_mean, running_var)
        else:
            scale, shift = RF.bn_eval_prepare(g, b, running_mean, running_var, eps)
            mean, invstd = running_mean, torch.rsqrt(running_var + eps)
        y, widx = RF.bn_relu_maxpool2d_fwd(xh, scale, shift, k, s, p)
        ctx.training, ctx.cfg = training, (tuple(xh.shape), k, s, p)
        ctx.save_for_backward(x2, widx, mean, invstd, gamma, beta, scale, shift)
        return nchw(y)

    @staticmethod
    def backward(ctx, dy):
        x2, widx, mean, invstd, gamma, beta, scale, shift = ctx.saved_tensors
        if not ctx.training:
            raise RuntimeError("BatchNorm2d backward in eval mode is not supported by replicann_b200")
        shape, k, s, p = ctx.cfg
        dz = RF.maxpool2d_bwd(nhwc(dy), None, shape, k, s, p, window_idx=widx)
        dg, acc = grad_target(gamma)
        db, acc_b = grad_target(beta)
        if acc != acc_b:
            raise RuntimeError("BatchNorm weight/bias gradient state out of sync")
        dx2, _ = RF.bn_bwd(dz.reshape(-1, shape[-1]), x2, None, mean, invstd, gamma.detach(), want_dres=False, dgamma=dg,
                           dbeta=db, accumulate=acc, fwd_scale=scale, fwd_shift=shift)
        _notify(gamma)
        _notify(beta)
        dx = nchw(dx2.reshape(shape)) if ctx.needs_input_grad[0] else None
        return (dx,) + (None,) * 10


class AvgPool2dFn(Function):
    @staticmethod
    def forward(ctx, x, k, s, p):
        xh = nhwc(x)
        ctx.cfg = (tuple(xh.shape), k, s, p)
        return nchw(RF.avgpool2d_fwd(xh, k, s, p))

    @staticmethod
    def backward(ctx, dy):
        shape, k, s, p = ctx.cfg
        return nchw(RF.avgpool2d_bwd(nhwc(dy), shape, k, s, p)), None, None, None


class GlobalAvgPoolFn(Function):
    @staticmethod
    def forward(ctx, x):
        xh = nhwc(x)
        ctx.shape = tuple(xh.shape)
        return nchw(RF.global_avgpool_fwd(xh))

    @staticmethod
    def backward(ctx, dy):
        return nchw(RF.global_avgpool_bwd(nhwc(dy), ctx.shape))


# ------------------------------------------------------------------------------------------ loss
class SoftmaxCrossEntropyFn(Function):
    @staticmethod
    def forward(ctx, logits, target):
        lg = logits if logits.is_contiguous() else logits.contiguous()
        loss, dl = RF.softmax_ce(lg, target, want_grad=logits.requires_grad)
        ctx.save_for_backward(dl)
        return loss

    @staticmethod
    def backward(ctx, dloss):
        (dl,) = ctx.saved_tensors
        ctx_one = getattr(SoftmaxCrossEntropyFn, "assume_unit_grad", False)
        return (dl if ctx_one else dl * dloss.to(dl.dtype)), None


class SoftmaxFn(Function):
    @staticmethod
    def forward(ctx, x):
        cols = x.shape[-1]
        x2 = x.reshape(-1, cols)
        y = RF.softmax_fwd(x2 if x2.is_contiguous() else x2.contiguous())
        ctx.save_for_backward(y)
        ctx.shape = x.shape
        return y.reshape(x.shape)

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        d2 = dy.reshape(y.shape)
        return RF.softmax_bwd(d2 if d2.is_contiguous() else d2.contiguous(), y).reshape(ctx.shape)


# ------------------------------------------------------------------------------------------ embeddings
class EmbeddingFn(Function):
    """out = tok[ids] (+ pos[:T]) -- torch.nn.Embedding lookups of a token and a learned positional table in one
    kernel; backward is the deterministic scatter-add / batch sum of rnn_embedding_bwd, written straight into the
    tables' .grad (a token table tied to the LM head then receives its second contribution by accumulation)."""

    @staticmethod
    def forward(ctx, ids, tok, pos):
        ids = ids if ids.is_contiguous() else ids.contiguous()
        T = ids.shape[1]
        ctx.save_for_backward(ids, tok, pos)
        return RF.embedding_fwd(ids, tok.detach(), pos.detach()[:T] if pos is not None else None)

    @staticmethod
    def backward(ctx, dy):
        ids, tok, pos = ctx.saved_tensors
        dy = dy if dy.is_contiguous() else dy.contiguous()
        T = ids.shape[1]
        dtok = dpos = None
        if ctx.needs_input_grad[1]:
            dtok = _write_param_grad(tok, lambda out, acc: RF.embedding_bwd(ids, dy, dtok=out, accumulate_tok=acc))
        if pos is not None and ctx.needs_input_grad[2]:
            def pg(out, acc):
                if not acc and out.shape[0] > T:
                    out[T:].zero_()          # positions beyond this batch's length get no gradient
                RF.embedding_bwd(ids, dy, dpos=out[:T], accumulate_pos=acc)
            dpos = _write_param_grad(pos, pg)
        return None, dtok, dpos


# ------------------------------------------------------------------------------------------ attention
class AttentionFn(Function):
    """z = dropout(softmax(q k^T * scale (+ mask))) v for all heads at once (nn/attention.py:28-50).

    `packed=True`: the first argument is the packed projection [B, T, 3*H*hs] (q | k | v column blocks) and
    a single packed gradient is returned; otherwise q, k, v are separate [B, T(k), H*hs] tensors."""

    @staticmethod
    def forward(ctx, q_or_qkv, k, v, mask, scale, p_drop, n_heads, hs, seed, offset, packed):
        Hh = n_heads * hs
        if packed:
            qkv = q_or_qkv if q_or_qkv.is_contiguous() else q_or_qkv.contiguous()
            q, k, v = qkv[..., :Hh], qkv[..., Hh:2 * Hh], qkv[..., 2 * Hh:]
        else:
            q = q_or_qkv if q_or_qkv.stride(-1) == 1 else q_or_qkv.contiguous()
            k = k if k.stride(-1) == 1 else k.contiguous()
            v = v if v.stride(-1) == 1 else v.contiguous()
        mask = RF.broadcast_mask(mask, q, k, v, n_heads, hs)
        # bf16 with a 64-multiple head size: contractions on the tensor cores (batched tcgen05 GEMMs over all heads);
        # anything else: the fused fp32-FMA kernel
        # fp32 large enough to pay for operand splits: the same pipeline at fp32 accuracy (three-piece bf16 splits)
        ctx.tc = 3 if RF.attention_fused_ok(q, k, v, n_heads, hs) else (
            1 if RF.attention_tc_ok(q, k, v, n_heads, hs) else (2 if RF.attention_f32_tc_ok(q, k, v, n_heads, hs) else 0))
        ctx.cfg = (scale, p_drop, n_heads, hs, seed, offset, packed)
        if ctx.tc == 3:
            # fused tcgen05 kernel: scores and probabilities stay on the SM; only z and the log-sum-exp are kept
            z, lse = RF.attention_fused_fwd(q, k, v, n_heads, hs, scale, mask, p_drop, seed, offset)
            ctx.save_for_backward(q, k, v, z, lse, mask)
            return z
        if ctx.tc:
            fwd = RF.attention_tc_fwd if ctx.tc == 1 else RF.attention_f32_tc_fwd
            z, saved = fwd(q, k, v, n_heads, hs, scale, mask, p_drop, seed, offset)
            ctx.save_for_backward(q, k, v, saved, mask)
            return z
        z, lse = RF.attention_fwd(q, k, v, n_heads, hs, scale, mask, p_drop, seed, offset)
        ctx.save_for_backward(q, k, v, z, lse, mask)
        return z

    @staticmethod
    def backward(ctx, dz):
        scale, p_drop, n_heads, hs, seed, offset, packed = ctx.cfg
        if ctx.tc in (1, 2):
            q, k, v, saved, mask = ctx.saved_tensors
        else:
            q, k, v, z, lse, mask = ctx.saved_tensors
        dz = dz if dz.stride(-1) == 1 else dz.contiguous()
        Hh = n_heads * hs
        if packed:
            dqkv = torch.empty((q.shape[0], q.shape[1], 3 * Hh), dtype=q.dtype, device=q.device)
            dq, dk, dv = dqkv[..., :Hh], dqkv[..., Hh:2 * Hh], dqkv[..., 2 * Hh:]
        else:
            dq = torch.empty_strided(q.shape, q.stride(), dtype=q.dtype, device=q.device)
            dk = torch.empty_strided(k.shape, k.stride(), dtype=k.dtype, device=k.device)
            dv = torch.empty_strided(v.shape, v.stride(), dtype=v.dtype, device=v.device)
        if ctx.tc == 3:
            if dz.stride(0) % 8 or dz.stride(1) % 8 or dz.data_ptr() % 16:
                dz = dz.contiguous()
            RF.attention_fused_bwd(q, k, v, z, dz, lse, n_heads, hs, scale, dq, dk, dv, mask, p_drop, seed, offset)
        elif ctx.tc:
            if dz.stride(0) % 8 or dz.stride(1) % 8 or dz.data_ptr() % 16:
                dz = dz.contiguous()
            bwd = RF.attention_tc_bwd if ctx.tc == 1 else RF.attention_f32_tc_bwd
            bwd(q, k, v, saved, dz, n_heads, hs, scale, dq, dk, dv, mask, p_drop, seed, offset)
        else:
            RF.attention_bwd(q, k, v, z, dz, lse, n_heads, hs, scale, dq, dk, dv, mask, p_drop, seed, offset)
        if packed:
            return dqkv, None, None, None, None, None, None, None, None, None, None
        return dq, dk, dv, None, None, None, None, None, None, None, None


# ------------------------------------------------------------------------------------------ fused residual block
class BottleneckFn(Function):
    """One ResNet bottleneck block as a single autograd node:
        y = relu(bn3(conv3(relu(bn2(conv2(relu(bn1(conv1(x)))))))) + identity),  identity = x | down_bn(down_conv(x))
    Forward runs the same kernels as the layer-by-layer path.  The hand-written backward chains them without
    autograd in between, which lets the input gradient be produced in ONE buffer: the residual-branch gradient
    written by the bn3 backward kernel is the accumulation target of the conv1 (and downsample) dgrad kernels,
    so the elementwise `grad_a + grad_b` pass autograd would otherwise run over the block input disappears.
    """

    @staticmethod
    def forward(ctx, x, stride, training, cfg, w1, g1, b1, w2, g2, b2, w3, g3, b3, wd, gd, bd, *stats):
        # stats: (rm1, rv1, rm2, rv2, rm3, rv3[, rmd, rvd]); cfg: ((eps, momentum) x 4)
        xh = nhwc(x)
        k1, k2, k3 = nhwc(w1), nhwc(w2), nhwc(w3)

        def conv_bn(xin, k, cs, cp, gamma, beta, rm, rv, em, residual, relu):
            """conv -> BatchNorm(+residual)(+ReLU).  In training the batch statistics come out of the convolution's
            own epilogue (no separate pass over the activation)."""
            g_, b_ = gamma.detach(), beta.detach()
            if training:
                y4, mean, invstd, scale, shift = RF.conv2d_fprop_bn_stats(xin, k, g_, b_, em[0], em[1], rm, rv,
                                                                          stride=cs, pad=cp)
            else:
                y4 = RF.conv2d_fprop(xin, k, stride=cs, pad=cp)
                scale, shift = RF.bn_eval_prepare(g_, b_, rm, rv, em[0])
                mean, invstd = rm, torch.rsqrt(rv + em[0])
            C_ = y4.shape[-1]
            y2 = y4.reshape(-1, C_)
            r2 = residual.reshape(-1, C_) if residual is not None else None
            if residual is not None and relu:
                z2, mask = RF.bn_apply(y2, scale, shift, r2, relu, want_mask=True)
            else:
                z2, mask = RF.bn_apply(y2, scale, shift, r2, relu), None
            return y4, z2.reshape(y4.shape), (mean, invstd, scale, shift), mask

        y1, z1, s1, _ = conv_bn(xh, k1, 1, 0, g1, b1, stats[0], stats[1], cfg[0], None, True)
        y2, z2, s2, _ = conv_bn(z1, k2, stride, 1, g2, b2, stats[2], stats[3], cfg[1], None, True)
        if wd is not None:
            yd, idt, sd, _ = conv_bn(xh, nhwc(wd), stride, 0, gd, bd, stats[6], stats[7], cfg[3], None, False)
        else:
            yd, idt, sd = None, xh, None
        y3, out, s3, omask = conv_bn(z2, k3, 1, 0, g3, b3, stats[4], stats[5], cfg[2], idt, True)
        ctx.stride, ctx.training, ctx.has_down = stride, training, wd is not None
        ctx.save_for_backward(xh, w1, g1, b1, w2, g2, b2, w3, g3, b3, wd, gd, bd, y1, z1, y2, z2, y3, yd, omask,
                              *s1, *s2, *s3, *(sd if sd is not None else ()))
        return nchw(out)

    @staticmethod
    def backward(ctx, dout):
        if not ctx.training:
            raise RuntimeError("BottleneckFn backward in eval mode is not supported")
        sv = ctx.saved_tensors
        (xh, w1, g1, b1, w2, g2, b2, w3, g3, b3, wd, gd, bd, y1, z1, y2, z2, y3, yd, omask) = sv[:20]
        s1, s2, s3 = sv[20:24], sv[24:28], sv[28:32]
        sd = sv[32:36] if ctx.has_down else None
        stride = ctx.stride
        k1, k2, k3 = nhwc(w1), nhwc(w2), nhwc(w3)
        dout_h = nhwc(dout)

        def bn_b(dy4, y4, st, gamma, beta, mask, want_dres, from_x):
            C_ = y4.shape[-1]
            dg, acc = grad_target(gamma)
            db, acc_b = grad_target(beta)
            dx2, dres2 = RF.bn_bwd(dy4.reshape(-1, C_), y4.reshape(-1, C_), None, st[0], st[1], gamma.detach(),
                                   want_dres=want_dres, dgamma=dg, dbeta=db, accumulate=acc or acc_b,
                                   fwd_scale=st[2] if from_x else None, fwd_shift=st[3] if from_x else None,
                                   relu_mask=mask)
            _notify(gamma); _notify(beta)
            return dx2.reshape(y4.shape), (dres2.reshape(y4.shape) if dres2 is not None else None)

        dev = dout.device
        use_side = RF.side_enabled()
        side_group: list = []   # (param, operands) of the weight-gradient kernels issued on the side stream

        def wgrad(w, xin, dy, s, p):
            kshape = tuple(nhwc(w).shape)
            if not (use_side and w.is_leaf):
                _write_param_grad(w, lambda o, acc: RF.conv2d_wgrad(xin, dy, kshape, stride=s, pad=p,
                                                                    out=nhwc_param_view(o), accumulate=acc))
                return
            # off the critical path: nothing in this backward pass reads dW, so the kernel goes to the side stream
            # where it overlaps the HBM-bound BatchNorm kernels of the main chain
            tgt, acc = grad_target(w)
            with RF.side_stream(dev):
                RF.conv2d_wgrad(xin, dy, kshape, stride=s, pad=p, out=nhwc_param_view(tgt), accumulate=acc)
            side_group.append((w, xin, dy, tgt))

        # bn3 (+ residual, ReLU): gradient of conv3's output and of the identity branch
        # Order of issue: each wgrad is enqueued AFTER the dgrad of the same layer, so on the side stream it starts
        # once that dgrad has finished and runs beside the BatchNorm backward kernels that follow on the main
        # stream (tensor-core work next to HBM streaming work) instead of contending with the dgrad for the SMs.
        dy3, dres = bn_b(dout_h, y3, s3, g3, b3, omask, True, False)
        dz2 = RF.conv2d_dgrad(dy3, k3, tuple(z2.shape))
        wgrad(w3, z2, dy3, 1, 0)
        dy2, _ = bn_b(dz2, y2, s2, g2, b2, None, False, True)
        dz1 = RF.conv2d_dgrad(dy2, k2, tuple(z1.shape), stride=stride, pad=1)
        wgrad(w2, z1, dy2, stride, 1)
        dy1, _ = bn_b(dz1, y1, s1, g1, b1, None, False, True)
        need_dx = ctx.needs_input_grad[0]
        dx = None
        if ctx.has_down:
            kd = nhwc(wd)
            dyd, _ = bn_b(dres, yd, sd, gd, bd, None, False, False)
            if need_dx:
                dx = RF.conv2d_dgrad(dyd, kd, tuple(xh.shape), stride=stride)
                RF.conv2d_dgrad(dy1, k1, tuple(xh.shape), out=dx, accumulate=True)
            wgrad(wd, xh, dyd, stride, 0)
        elif need_dx:
            # identity shortcut: dres already IS d(out)/d(x) through the skip; conv1's dgrad accumulates into it
            dx = RF.conv2d_dgrad(dy1, k1, tuple(xh.shape), out=dres, accumulate=True)
        wgrad(w1, xh, dy1, 1, 0)
        if side_group:
            params = [g[0] for g in side_group]
            RF.side_commit(dev, side_group, lambda: [_notify(p_) for p_ in params])
            # join with a lag of one block: this block's weight gradients overlap the next block's main chain
            RF.side_join(dev, keep=1)
            RF.side_join_at_end_of_backward(dev)
        return (nchw(dx) if dx is not None else None,) + (None,) * (15 + (8 if ctx.has_down else 6))
