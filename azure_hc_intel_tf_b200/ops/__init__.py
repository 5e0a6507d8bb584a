"""Layer-level ops used by the model zoo.

Every function takes the layer module (for its parameters and static geometry) and dispatches on
``module.cfg.kernels``:

* ``'b200'``  → ops/b200.py: hand-written sm_100a kernels from libb200cnn.so (NHWC bf16).
* ``'cudnn'`` → stock PyTorch ops (cuDNN/cuBLAS on GPU, native kernels on CPU). This is the measured
  baseline and the CPU plumbing path; it is never the product path on a GPU.

Counterparts in the reference stack are the MKL-DNN primitives TensorFlow calls when the launcher
passes ``--mkl=TRUE --data_format=NCHW`` (benchmark-scripts/run-tf-sing-ucx-openmpi.sh:72,76).
"""
from __future__ import annotations

import torch
import torch.nn.functional as F


def _b200():
    from . import b200
    return b200


def _vec8(c: int) -> bool:
    """The sm_100a kernels move 8-channel (16-byte) vectors; layers whose channel count is not a multiple of 8
    (DenseNet growth 12, CIFAR stems) take the stock ops instead."""
    return c % 8 == 0


def _sym(pad):
    (pt, pb), (pl, pr) = pad
    return pt == pb and pl == pr


# ------------------------------------------------------------------------------------ conv block

def _act(y, relu):
    """``relu``: 0 / False = none, 1 / True = ReLU, 2 = ReLU6 (MobileNet)."""
    if int(relu) == 2:
        return F.relu6(y, inplace=True)
    return F.relu(y, inplace=True) if relu else y


def conv_block(x, m, relu, residual=None):
    """conv (+bias) (+training-mode batch-norm (+residual)) (+ReLU | ReLU6); grouped when ``m.groups`` > 1."""
    if m.cfg.kernels == "b200" and _vec8(m.cout) and (_vec8(m.cin) or m.cin < 8):
        return _b200().conv_block(x, m, relu, residual=residual)
    (pt, pb), (pl, pr) = m.pad
    w = m.weight if m.weight.dtype == x.dtype else m.weight.to(x.dtype)
    bias = None if (m.bn is not None or m.bias is None) else m.bias.to(x.dtype)
    groups = getattr(m, "groups", 1)
    if _sym(m.pad):
        y = F.conv2d(x, w, bias, m.stride, (pt, pl), 1, groups)
    else:
        y = F.conv2d(F.pad(x, (pl, pr, pt, pb)), w, bias, m.stride, 0, 1, groups)
    if m.bn is not None:
        return batch_norm(y, m.bn, relu=relu, residual=residual)
    if residual is not None:
        y = y + residual
    return _act(y, relu)


def batch_norm(y, bn, relu=False, residual=None):
    """Training-mode BN; optional fused residual add and ReLU / ReLU6 (y = act(bn(y) + residual))."""
    if bn.cfg.kernels == "b200" and _vec8(y.shape[1]) and y.dim() == 4:
        return _b200().batch_norm(y, bn, relu, residual)
    out = F.batch_norm(y, bn.moving_mean, bn.moving_variance, bn.gamma, bn.beta, bn.training,
                       1.0 - bn.decay, bn.eps)
    if residual is not None:
        out = out + residual
    return _act(out, relu)


def conv_bn_add_relu(x, m, residual):
    """Last conv of a residual block: relu(bn(conv(x)) + residual)."""
    if m.cfg.kernels == "b200":
        return _b200().conv_block(x, m, True, residual=residual)
    (pt, pb), (pl, pr) = m.pad
    y = F.conv2d(x if _sym(m.pad) else F.pad(x, (pl, pr, pt, pb)), m.weight, None, m.stride,
                 (pt, pl) if _sym(m.pad) else 0)
    return batch_norm(y, m.bn, relu=True, residual=residual)


def bottleneck(x, blk):
    """ResNet bottleneck block: relu(bn3(conv3(relu(bn2(conv2(relu(bn1(conv1(x)))))))) + shortcut(x)). On the b200
    kernels the block is one autograd node (ops/b200.py ``_Bottleneck``) so the gradient sum at the block input is
    fused with the output ReLU's gate; B200_FUSED_BLOCK=0 falls back to chaining the per-layer nodes."""
    if blk.cfg.kernels == "b200" and _fused_block() and _b200().bottleneck_supported(blk):
        return _b200().bottleneck(x, blk)
    sc = x if blk.shortcut is None else blk.shortcut(x)
    y = blk.conv2(blk.conv1(x))
    return conv_bn_add_relu(y, blk.conv3, sc)


def _fused_block() -> bool:
    import os
    return os.environ.get("B200_FUSED_BLOCK", "1") != "0"


# ------------------------------------------------------------------------------------ dense / pool

def affine(x, m, relu: bool):
    if m.cfg.kernels == "b200":
        return _b200().affine(x, m, relu)
    y = F.linear(x, m.weight, m.bias)
    return F.relu(y, inplace=True) if relu else y


def pool(x, m):
    if m.cfg.kernels == "b200" and _vec8(x.shape[1]):
        return _b200().pool(x, m)
    (pt, pb), (pl, pr) = m.pad
    if m.kind == "max":
        if not _sym(m.pad) or pt or pl:
            x = F.pad(x, (pl, pr, pt, pb), value=float("-inf"))
        return F.max_pool2d(x, m.k, m.stride)
    # TF 'SAME' average pooling divides by the number of valid (unpadded) taps.
    if _sym(m.pad):
        return F.avg_pool2d(x, m.k, m.stride, (pt, pl), count_include_pad=False)
    ones = torch.ones_like(x[:1, :1])
    num = F.avg_pool2d(F.pad(x, (pl, pr, pt, pb)), m.k, m.stride)
    den = F.avg_pool2d(F.pad(ones, (pl, pr, pt, pb)), m.k, m.stride)
    return num / den


def spatial_mean(x, cfg):
    if cfg.kernels == "b200" and _vec8(x.shape[1]):
        return _b200().spatial_mean(x)
    return x.mean(dim=(2, 3))


def add_relu(a, b, cfg):
    if cfg.kernels == "b200":
        return _b200().add_relu(a, b)
    return F.relu(a + b, inplace=True)


def relu(x, cfg):
    """Out-of-place ReLU (the input usually feeds other branches as well)."""
    if cfg.kernels == "b200" and x.dim() == 4 and _vec8(x.shape[1]):
        return _b200().relu(x)
    return F.relu(x)


def add(a, b, cfg):
    """Residual sum without activation (pre-activation ResNets)."""
    if cfg.kernels == "b200":
        return _b200().add_act(a, b, False)
    return a + b


def concat_channels(xs):
    # channels_last tensors: the concat is a strided copy into the channel-interleaved output.
    return torch.cat(xs, dim=1)


# ------------------------------------------------------------------------------------ loss

def softmax_cross_entropy(logits, labels, cfg=None):
    """Mean softmax cross-entropy over the batch (tf_cnn_benchmarks ``loss_function``)."""
    if cfg is not None and cfg.kernels == "b200":
        return _b200().softmax_cross_entropy(logits, labels)
    return F.cross_entropy(logits.float(), labels)


def l2_loss(params, weight_decay: float):
    """tf_cnn_benchmarks adds ``weight_decay * sum(tf.nn.l2_loss(v))`` = wd/2·Σ‖v‖² over trainable
    variables. Its gradient, ``wd·v``, is folded into the optimizer kernels; this function only reports
    the scalar for the ``total_loss`` column."""
    tot = None
    for p in params:
        s = p.detach().float().pow(2).sum()
        tot = s if tot is None else tot + s
    return 0.5 * weight_decay * tot if tot is not None else torch.zeros(())
