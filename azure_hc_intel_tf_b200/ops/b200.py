"""Layer ops on the hand-written sm_100a kernels (libb200cnn.so), NHWC bf16.

Each public function mirrors one in ops/__init__.py. Tensors are logical-NCHW ``channels_last`` (so the
storage is NHWC and ``data_ptr()`` can be handed to the kernels unchanged).

Gradient hand-off: when a parameter's ``.grad`` is a view into a fusion buffer (the optimizer attached
``grad_ready``), backward writes the fp32 gradient straight into that view and calls ``grad_ready()`` —
the Horovod "enqueue tensor" moment — instead of returning it to autograd. Otherwise the gradient is
returned normally.

Counterpart in the reference stack: TensorFlow + MKL-DNN primitives under ``--mkl=TRUE``
(benchmark-scripts/run-tf-sing-ucx-openmpi.sh:76); SURVEY.md §2.6 K1–K9.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from .. import _native
from .._cuda_sigs import PoolGeom
from . import conv_impl

_BF16 = torch.bfloat16
_scratch = {}
launch_count = 0          # kernels of ours enqueued since the last reset (bench.py reports it)


def lib():
    return _native.cuda()


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _check(rc, name):
    global launch_count
    launch_count += 1
    if rc != 0:
        raise RuntimeError(f"{name} failed ({rc}): {lib().b200_last_error().decode()}")


def count_launch(n=1):
    global launch_count
    launch_count += n


def _ptr(t: Optional[torch.Tensor]):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def scratch(device) -> torch.Tensor:
    """Per-device reduction scratch (partials + ticket counters). Kernels using it run on one stream."""
    key = (device.index, torch.cuda.current_stream().cuda_stream)
    s = _scratch.get(key)
    if s is None:
        s = torch.zeros(8 << 20, dtype=torch.uint8, device=device)
        _scratch[key] = s
    return s


def nhwc(x: torch.Tensor) -> torch.Tensor:
    """Ensure bf16 channels_last storage; returns the same logical-NCHW tensor."""
    if x.dtype != _BF16:
        x = x.to(_BF16)
    if x.dim() == 4 and not x.is_contiguous(memory_format=torch.channels_last):
        x = x.contiguous(memory_format=torch.channels_last)
    return x


def empty_nhwc(n, c, h, w, device, dtype=_BF16):
    return torch.empty((n, h, w, c), dtype=dtype, device=device).permute(0, 3, 1, 2)


def _direct(p: Optional[torch.Tensor]):
    """The fp32 fusion-buffer view to write this parameter's gradient into, or None."""
    if p is not None and p.requires_grad and getattr(p, "grad_ready", None) is not None and p.grad is not None:
        return p.grad
    return None


def _deliver(p, grad_view, value):
    """Return value for autograd: None if the gradient was written in place (and signal readiness)."""
    if grad_view is not None:
        p.grad_ready()
        return None
    return value


def lowp(p: torch.Tensor) -> torch.Tensor:
    lp = getattr(p, "lp", None)
    return lp if lp is not None else p.detach().to(_BF16)


# ------------------------------------------------------------------------------------ BN pieces

def bn_forward(y, bn, residual, relu, sums=None, want_mask=False):
    """y: conv output (bf16 NHWC). Returns (z, saved) with saved = (mean, rstd, scale, shift). ``sums``
    ([rows,2,C] fp32 partial Σy, Σy²) comes from the conv epilogue when available, which saves one read of y."""
    n, c, h, w = y.shape
    rows = n * h * w
    dev = y.device
    stats = torch.empty((4, c), dtype=torch.float32, device=dev)
    mean, rstd, scale, shift = stats[0], stats[1], stats[2], stats[3]
    training = bn.training
    if training and sums is not None:
        _check(lib().b200_bn_finalize_partials(_ptr(sums), sums.shape[0], rows, c, _ptr(bn.gamma), _ptr(bn.beta),
                                               _ptr(bn.moving_mean), _ptr(bn.moving_variance), bn.decay, bn.eps,
                                               _ptr(mean), _ptr(rstd), _ptr(scale), _ptr(shift), _stream()),
               "bn_finalize_partials")
    elif training:
        _check(lib().b200_bn_stats(_ptr(y), rows, c, _ptr(bn.gamma), _ptr(bn.beta), _ptr(bn.moving_mean),
                                   _ptr(bn.moving_variance), bn.decay, bn.eps, _ptr(mean), _ptr(rstd), _ptr(scale),
                                   _ptr(shift), _ptr(scratch(dev)), _stream()), "bn_stats")
    else:
        rstd.copy_(torch.rsqrt(bn.moving_variance + bn.eps))
        mean.copy_(bn.moving_mean)
        scale.copy_(bn.gamma.detach() * rstd)
        shift.copy_(bn.beta.detach() - mean * scale)
    z = empty_nhwc(n, c, h, w, dev)
    # with a residual the ReLU mask cannot be recomputed from y: keep it as one bit per element for the backward;
    # ReLU6 (relu == 2) always keeps it - its pass-through set is 0 < f < 6, which the backward kernels take from the mask
    keep = relu and want_mask and (residual is not None or int(relu) == 2)
    mask = torch.empty(rows * c // 8, dtype=torch.uint8, device=dev) if keep else None
    _check(lib().b200_bn_apply(_ptr(y), _ptr(residual), _ptr(z), _ptr(mask), rows, c, _ptr(scale), _ptr(shift), int(relu),
                               _stream()), "bn_apply")
    return (z, stats, mask) if want_mask else (z, stats)


def bn_backward(dz, y, z_for_mask, stats, relu, want_dres, dgamma_out, dbeta_out, mask=None):
    n, c, h, w = y.shape
    rows = n * h * w
    dev = y.device
    mean, rstd, scale, shift = stats[0], stats[1], stats[2], stats[3]
    dgamma = dgamma_out if dgamma_out is not None else torch.empty(c, dtype=torch.float32, device=dev)
    dbeta = dbeta_out if dbeta_out is not None else torch.empty(c, dtype=torch.float32, device=dev)
    _check(lib().b200_bn_bwd_reduce(_ptr(dz), _ptr(y), _ptr(z_for_mask), _ptr(mask), rows, c, _ptr(mean), _ptr(rstd), _ptr(scale),
                                    _ptr(shift), int(relu), _ptr(dgamma), _ptr(dbeta), _ptr(scratch(dev)), _stream()),
           "bn_bwd_reduce")
    dy = empty_nhwc(n, c, h, w, dev)
    dres = empty_nhwc(n, c, h, w, dev) if want_dres else None
    _check(lib().b200_bn_bwd_apply(_ptr(dz), _ptr(y), _ptr(z_for_mask), _ptr(mask), _ptr(dy), _ptr(dres), rows, c, _ptr(mean),
                                   _ptr(rstd), _ptr(scale), _ptr(shift), _ptr(dgamma), _ptr(dbeta), int(relu),
                                   _stream()), "bn_bwd_apply")
    return dy, dres, dgamma, dbeta


# ------------------------------------------------------------------------------------ conv block

def _cb_forward(x, weight, m, relu, residual):
    """conv (+BN (+residual) (+ReLU) | +bias (+ReLU)) forward. Returns (z, saved) where ``saved`` is what
    ``_cb_backward`` needs: (x or the first layer's patch matrix, bf16 weight, y, mask-or-z, stats, first)."""
    x = nhwc(x)
    if residual is not None:
        residual = nhwc(residual)
    w_lp = lowp(weight)
    first = conv_impl.first_layer(x, m)       # Cin = 3: explicit patch matrix, kept for the weight gradient
    if first:
        y, sums, patches = conv_impl.first_fprop(x, w_lp, m, m.bn is not None and m.bn.training)
        x = patches
    if m.bn is not None:
        if not first:
            y, sums = conv_impl.fprop(x, w_lp, m, want_stats=m.bn.training)
        z, stats, mask = bn_forward(y, m.bn, residual, relu, sums, want_mask=True)
        # the ReLU mask is recomputed from y when there is no residual, else kept as one bit per element
        return z, (x, w_lp, y, mask, stats, first)
    if not first:
        y = conv_impl.fprop(x, w_lp, m)
    if m.bias is not None or relu:
        assert int(relu) != 2, "ReLU6 is only implemented behind batch-norm (MobileNet)"
        n, c, h, w = y.shape
        _check(lib().b200_bias_act(_ptr(y), _ptr(m.bias), _ptr(y), n * h * w, c, int(relu), _stream()), "bias_act")
    return y, (x, w_lp, None, y if relu else None, None, first)


def _cb_backward(saved, dz, m, relu, weight, gamma, beta, bias, need_dx, want_dres, gate=None):
    """Backward of ``_cb_forward``. ``gate``: ReLU bit mask of a LATER layer that ``dz`` has not passed yet (the
    projection shortcut of a residual block receives the block-output gradient before its ReLU gate); only for
    BN layers without their own ReLU. Returns (dx, dres, gw, ggamma, gbeta, gbias) with the parameter entries
    already run through ``_deliver`` (None when written into the fusion buffer)."""
    x, w_lp, y, z, stats, first = saved
    dz = nhwc(dz)
    dres = dgamma = dbeta = dbias = None
    g_gamma = g_beta = g_bias = None
    if m.bn is not None:
        g_gamma = _direct(gamma) if (gamma is not None and gamma.requires_grad) else None
        g_beta = _direct(beta)
        if gate is not None:
            assert not relu, "a gated gradient can only enter a layer without its own ReLU"
            dy, dres, dgamma, dbeta = bn_backward(dz, y, None, stats, True, want_dres, g_gamma, g_beta, mask=gate)
        else:
            dy, dres, dgamma, dbeta = bn_backward(dz, y, None, stats, relu, want_dres, g_gamma, g_beta, mask=z)
    else:
        assert gate is None
        n, c, h, w = dz.shape
        if bias is not None:
            g_bias = _direct(bias)
            dbias = g_bias if g_bias is not None else torch.empty(c, dtype=torch.float32, device=dz.device)
            dy = empty_nhwc(n, c, h, w, dz.device) if relu else dz
            _check(lib().b200_bias_act_bwd(_ptr(dz), _ptr(z), _ptr(dy) if relu else _ptr(None), _ptr(dbias),
                                           n * h * w, c, int(relu), _ptr(scratch(dz.device)), _stream()),
                   "bias_act_bwd")
        elif relu:
            dy = torch.where(z > 0, dz, torch.zeros_like(dz))
        else:
            dy = dz
    g_w = _direct(weight)
    if first:
        dw = conv_impl.first_wgrad(x, dy, m, out=g_w)      # x holds the saved patch matrix
        dx = None
    else:
        dw = conv_impl.wgrad(x, dy, m, out=g_w)
        dx = conv_impl.dgrad(dy, w_lp, m, x.shape) if need_dx else None
    return (dx, dres, _deliver(weight, g_w, dw),
            _deliver(gamma, g_gamma, dgamma) if (gamma is not None and gamma.requires_grad) else None,
            _deliver(beta, g_beta, dbeta) if beta is not None else None,
            _deliver(bias, g_bias, dbias) if bias is not None else None)


class _ConvBlock(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, gamma, beta, bias, residual, m, relu):
        z, saved = _cb_forward(x, weight, m, relu, residual)
        ctx.m, ctx.relu, ctx.has_res, ctx.first = m, relu, residual is not None, saved[5]
        ctx.save_for_backward(*saved[:5])
        ctx.weight, ctx.gamma, ctx.beta, ctx.bias = weight, gamma, beta, bias
        return z

    @staticmethod
    def backward(ctx, dz):
        saved = tuple(ctx.saved_tensors) + (ctx.first,)
        dx, dres, gw, gg, gb, gbias = _cb_backward(saved, dz, ctx.m, ctx.relu, ctx.weight, ctx.gamma, ctx.beta, ctx.bias,
                                                   ctx.needs_input_grad[0], ctx.has_res)
        return dx, gw, gg, gb, gbias, dres, None, None


class _Bottleneck(torch.autograd.Function):
    """A whole ResNet bottleneck block (conv1-bn-relu, conv2-bn-relu, conv3-bn, [projection conv-bn], add, relu) as
    ONE autograd node. What that buys over chaining ``_ConvBlock`` nodes: the two gradient branches that meet at the
    block input are summed here with one kernel that also applies the output ReLU's gate to the shortcut branch
    (``b200_add_masked``), so the gated copy of the block-output gradient is never written (one activation-sized
    write less per block) and autograd's own accumulation add disappears from the step.

    Inputs: x, then (weight, gamma, beta) of conv1, conv2, conv3 and - for a projection shortcut - of the
    shortcut conv. The reference runs this block as separate TF ops fused by MKL-DNN where it can
    (tf_cnn_benchmarks resnet_model bottleneck_block, selected by run-tf-sing-ucx-openmpi.sh:34)."""

    @staticmethod
    def forward(ctx, x, blk, *params):
        convs = [blk.conv1, blk.conv2, blk.conv3] + ([blk.shortcut] if blk.shortcut is not None else [])
        ctx.blk, ctx.params, ctx.nconv = blk, params, len(convs)
        x = nhwc(x)
        tensors, firsts = [], []
        sc = x
        if blk.shortcut is not None:
            sc, sv = _cb_forward(x, params[9], blk.shortcut, False, None)
            tensors += sv[:5]
            firsts.append(sv[5])
        h = x
        saved3 = []
        for i, m in enumerate(convs[:3]):
            h, sv = _cb_forward(h, params[3 * i], m, True, sc if i == 2 else None)
            saved3.append(sv)
        for sv in saved3:
            tensors += sv[:5]
        ctx.save_for_backward(*tensors)
        return h

    @staticmethod
    def backward(ctx, dz):
        blk, params = ctx.blk, ctx.params
        t = list(ctx.saved_tensors)
        proj = blk.shortcut is not None
        off = 5 if proj else 0
        sv = [tuple(t[off + 5 * i: off + 5 * i + 5]) + (False,) for i in range(3)]
        dz = nhwc(dz)
        grads = [None] * len(params)
        # main path, last conv first; conv3's residual gradient stays implicit as (dz, its ReLU mask)
        g = dz
        for i in (2, 1, 0):
            m = (blk.conv1, blk.conv2, blk.conv3)[i]
            g, _, gw, gg, gb, _ = _cb_backward(sv[i], g, m, True, params[3 * i], params[3 * i + 1], params[3 * i + 2],
                                              None, True, False)
            grads[3 * i: 3 * i + 3] = [gw, gg, gb]
        mask3 = sv[2][3]
        n, c, hh, ww = (g.shape if g is not None else (0, 0, 0, 0))
        dx = None
        if ctx.needs_input_grad[0]:
            dx = empty_nhwc(n, c, hh, ww, g.device)
            if proj:
                svp = tuple(t[0:5]) + (False,)
                gp, _, gw, gg, gb, _ = _cb_backward(svp, dz, blk.shortcut, False, params[9], params[10], params[11], None,
                                                    True, False, gate=mask3)
                grads[9:12] = [gw, gg, gb]
                _check(lib().b200_add(_ptr(g), _ptr(gp), _ptr(dx), g.numel(), _stream()), "add")
            else:
                _check(lib().b200_add_masked(_ptr(g), _ptr(dz), _ptr(mask3), _ptr(dx), g.numel(), _stream()), "add_masked")
        elif proj:
            svp = tuple(t[0:5]) + (False,)
            _, _, gw, gg, gb, _ = _cb_backward(svp, dz, blk.shortcut, False, params[9], params[10], params[11], None,
                                               False, False, gate=mask3)
            grads[9:12] = [gw, gg, gb]
        return (dx, None) + tuple(grads)


def bottleneck_supported(blk) -> bool:
    """The fused node covers the standard block: three conv+BN layers without bias, identity or conv+BN shortcut."""
    convs = [blk.conv1, blk.conv2, blk.conv3]
    sc = blk.shortcut
    if sc is not None and not (hasattr(sc, "weight") and getattr(sc, "bn", None) is not None):
        return False
    if sc is not None:
        convs.append(sc)
    return all(m.bn is not None and m.bias is None and m.bn.gamma is not None and m.bn.beta is not None for m in convs)


def bottleneck(x, blk):
    params = []
    for m in [blk.conv1, blk.conv2, blk.conv3] + ([blk.shortcut] if blk.shortcut is not None else []):
        params += [m.weight, m.bn.gamma, m.bn.beta]
    return _Bottleneck.apply(x, blk, *params)


def conv_block(x, m, relu, residual=None):
    bn = m.bn
    return _ConvBlock.apply(x, m.weight, bn.gamma if bn is not None else None, bn.beta if bn is not None else None,
                            m.bias, residual, m, relu)


class _BatchNorm(torch.autograd.Function):
    """Standalone BatchNorm(+residual, +ReLU) for graphs that do not reach it through conv_block."""

    @staticmethod
    def forward(ctx, y, gamma, beta, residual, bn, relu):
        y = nhwc(y)
        if residual is not None:
            residual = nhwc(residual)
        z, stats, mask = bn_forward(y, bn, residual, relu, None, want_mask=True)
        ctx.save_for_backward(y, mask, stats)
        ctx.bn, ctx.relu, ctx.has_res, ctx.gamma, ctx.beta = bn, relu, residual is not None, gamma, beta
        return z

    @staticmethod
    def backward(ctx, dz):
        y, mask, stats = ctx.saved_tensors
        gamma, beta = ctx.gamma, ctx.beta
        g_gamma = _direct(gamma) if gamma.requires_grad else None
        g_beta = _direct(beta)
        dy, dres, dgamma, dbeta = bn_backward(nhwc(dz), y, None, stats, ctx.relu, ctx.has_res, g_gamma, g_beta, mask=mask)
        return (dy, _deliver(gamma, g_gamma, dgamma) if gamma.requires_grad else None, _deliver(beta, g_beta, dbeta),
                dres, None, None)


def batch_norm(y, bn, relu, residual):
    return _BatchNorm.apply(y, bn.gamma, bn.beta, residual, bn, relu)


# ------------------------------------------------------------------------------------ affine

class _Affine(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, m, relu):
        x2 = x.to(_BF16).contiguous()
        w_lp = lowp(weight)
        head = (m.cout % 8 != 0)
        y = conv_impl.gemm_nt(x2, w_lp, out_fp32=head)      # [B, out]; the 1001-way head keeps fp32 logits
        b, c = y.shape
        if c % 8 == 0:
            _check(lib().b200_bias_act(_ptr(y), _ptr(bias), _ptr(y), b, c, int(relu), _stream()), "bias_act")
        else:                                                # 1001-way head: tiny, unaligned
            y = y + bias.to(y.dtype)
            if relu:
                y = torch.relu_(y)
        ctx.save_for_backward(x2, w_lp, y if relu else None)
        ctx.weight, ctx.bias, ctx.relu, ctx.out_features = weight, bias, relu, m.cout
        return y

    @staticmethod
    def backward(ctx, dy):
        x2, w_lp, y = ctx.saved_tensors
        weight, bias = ctx.weight, ctx.bias
        if dy.dtype != _BF16:
            dy = dy.to(_BF16)
        b, c = dy.shape
        g_b = _direct(bias)
        if c % 8 == 0:
            dy = dy.contiguous()
            dbias = g_b if g_b is not None else torch.empty(c, dtype=torch.float32, device=dy.device)
            d2 = torch.empty_like(dy) if ctx.relu else dy
            _check(lib().b200_bias_act_bwd(_ptr(dy), _ptr(y), _ptr(d2) if ctx.relu else _ptr(None), _ptr(dbias), b, c,
                                           int(ctx.relu), _ptr(scratch(dy.device)), _stream()), "bias_act_bwd")
            dy = d2
        else:
            # 1001-way head: tiny and unaligned; pad the columns to a multiple of 8 so TMA can address the rows
            if ctx.relu:
                dy = torch.where(y > 0, dy, torch.zeros_like(dy))
            dbias = dy.float().sum(0)
            if g_b is not None:
                g_b.copy_(dbias)
            ld = (c + 7) // 8 * 8
            dy_pad = dy.new_zeros((b, ld))
            dy_pad[:, :c] = dy
            dy = dy_pad
        g_w = _direct(weight)
        dw = conv_impl.gemm_tn_f32(dy, x2, out=g_w, out_features=ctx.out_features)   # [out, in] fp32
        dx = conv_impl.gemm_nn(dy, w_lp) if ctx.needs_input_grad[0] else None
        return dx, _deliver(weight, g_w, dw), _deliver(bias, g_b, dbias), None, None


def affine(x, m, relu):
    return _Affine.apply(x, m.weight, m.bias, m, relu)


# ------------------------------------------------------------------------------------ pooling

def _geom(x, m) -> PoolGeom:
    n, c, h, w = x.shape
    p, q = m.out_hw
    return PoolGeom(n, h, w, c, p, q, m.k[0], m.k[1], m.stride[0], m.stride[1], m.pad[0][0], m.pad[1][0])


class _Pool(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, m):
        x = nhwc(x)
        g = _geom(x, m)
        y = empty_nhwc(g.n, g.c, g.p, g.q, x.device)
        if m.kind == "max":
            idx = torch.empty((g.n, g.p, g.q, g.c), dtype=torch.uint8, device=x.device)
            _check(lib().b200_maxpool_fwd(_ptr(x), _ptr(y), _ptr(idx), C.byref(g), _stream()), "maxpool_fwd")
            ctx.save_for_backward(idx)
        else:
            _check(lib().b200_avgpool_fwd(_ptr(x), _ptr(y), C.byref(g), _stream()), "avgpool_fwd")
        ctx.g, ctx.kind = g, m.kind
        return y

    @staticmethod
    def backward(ctx, dy):
        g = ctx.g
        dy = nhwc(dy)
        dx = empty_nhwc(g.n, g.c, g.h, g.w, dy.device)
        if ctx.kind == "max":
            (idx,) = ctx.saved_tensors
            _check(lib().b200_maxpool_bwd(_ptr(dy), _ptr(idx), _ptr(dx), C.byref(g), _stream()), "maxpool_bwd")
        else:
            _check(lib().b200_avgpool_bwd(_ptr(dy), _ptr(dx), C.byref(g), _stream()), "avgpool_bwd")
        return dx, None


def pool(x, m):
    return _Pool.apply(x, m)


class _SpatialMean(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        x = nhwc(x)
        n, c, h, w = x.shape
        y = torch.empty((n, c), dtype=_BF16, device=x.device)
        _check(lib().b200_spatial_mean_fwd(_ptr(x), _ptr(y), n, h * w, c, _stream()), "spatial_mean_fwd")
        ctx.shape = (n, c, h, w)
        return y

    @staticmethod
    def backward(ctx, dy):
        n, c, h, w = ctx.shape
        dy = dy.to(_BF16).contiguous()
        dx = empty_nhwc(n, c, h, w, dy.device)
        _check(lib().b200_spatial_mean_bwd(_ptr(dy), _ptr(dx), n, h * w, c, _stream()), "spatial_mean_bwd")
        return dx


def spatial_mean(x):
    return _SpatialMean.apply(x)


class _AddAct(torch.autograd.Function):
    """act(a + b) outside a conv block: the residual sum of the CIFAR ResNets (ReLU) and of the pre-activation models
    (no activation). ResNet v1's sum rides in bn_apply (conv_block(residual=...)); this is for blocks whose sum is not
    behind a batch-norm. Reference role: the Add / Relu ops MKL-DNN fuses under --mkl=TRUE (run-tf-sing-ucx-openmpi.sh:76)."""

    @staticmethod
    def forward(ctx, a, b, relu):
        a, b = nhwc(a), nhwc(b)
        n, c, h, w = a.shape
        z = empty_nhwc(n, c, h, w, a.device)
        _check(lib().b200_add_act(_ptr(a), _ptr(b), _ptr(z), a.numel(), int(relu), _stream()), "add_act")
        ctx.relu = relu
        if relu:
            ctx.save_for_backward(z)
        return z

    @staticmethod
    def backward(ctx, dz):
        if not ctx.relu:
            return dz, dz, None
        (z,) = ctx.saved_tensors
        dz = nhwc(dz)
        n, c, h, w = dz.shape
        dx = empty_nhwc(n, c, h, w, dz.device)
        _check(lib().b200_relu_bwd(_ptr(dz), _ptr(z), _ptr(dx), dz.numel(), _stream()), "relu_bwd")
        return dx, dx, None


class _Relu(torch.autograd.Function):
    """Standalone ReLU (the pre-activation in front of NASNet's separable convolutions, whose input also feeds other
    branches un-activated): bias_act_kernel without a bias forward, relu_bwd_kernel backward (gate from the output)."""

    @staticmethod
    def forward(ctx, x):
        x = nhwc(x)
        n, c, h, w = x.shape
        z = empty_nhwc(n, c, h, w, x.device)
        _check(lib().b200_bias_act(_ptr(x), _ptr(None), _ptr(z), n * h * w, c, 1, _stream()), "bias_act")
        ctx.save_for_backward(z)
        return z

    @staticmethod
    def backward(ctx, dz):
        (z,) = ctx.saved_tensors
        dz = nhwc(dz)
        n, c, h, w = dz.shape
        dx = empty_nhwc(n, c, h, w, dz.device)
        _check(lib().b200_relu_bwd(_ptr(dz), _ptr(z), _ptr(dx), dz.numel(), _stream()), "relu_bwd")
        return dx


def relu(x):
    return _Relu.apply(x)


def add_act(a, b, relu):
    if a.dim() != 4 or a.shape != b.shape or a.numel() % 8 or a.shape[1] % 8:
        return torch.relu_(a + b) if relu else a + b          # shapes the 8-channel vectors do not cover
    return _AddAct.apply(a, b, bool(relu))


def add_relu(a, b):
    return add_act(a, b, True)


# ------------------------------------------------------------------------------------ loss / input

class _SoftmaxXent(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels):
        lg = logits.float().contiguous()
        b, classes = lg.shape
        ld_d = (classes + 7) // 8 * 8
        rows = torch.empty(b, dtype=torch.float32, device=lg.device)
        loss = torch.empty((), dtype=torch.float32, device=lg.device)
        d = torch.empty((b, ld_d), dtype=_BF16, device=lg.device)
        _check(lib().b200_softmax_xent(_ptr(lg), classes, _ptr(labels), b, classes, _ptr(rows), _ptr(loss), _ptr(d),
                                       ld_d, 1.0, _stream()), "softmax_xent")
        count_launch(1)
        ctx.save_for_backward(d)
        ctx.classes, ctx.dtype = classes, logits.dtype
        return loss

    @staticmethod
    def backward(ctx, g):
        (d,) = ctx.saved_tensors
        out = d[:, :ctx.classes]
        # g is 1.0 in training (loss.backward()); honour other scalings without a sync
        return (out * g.to(out.dtype)).to(ctx.dtype), None


def softmax_cross_entropy(logits, labels):
    return _SoftmaxXent.apply(logits, labels)


def normalize_u8(u8_nhwc: torch.Tensor, c_out: Optional[int] = None) -> torch.Tensor:
    n, h, w, c = u8_nhwc.shape
    c_out = c_out or c
    y = torch.empty((n, h, w, c_out), dtype=_BF16, device=u8_nhwc.device)
    _check(lib().b200_normalize_u8(_ptr(u8_nhwc), _ptr(y), n * h * w, c, c_out, _stream()), "normalize_u8")
    return y.permute(0, 3, 1, 2)
