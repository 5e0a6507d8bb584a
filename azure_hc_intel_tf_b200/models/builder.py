"""ConvNetBuilder-style layer factory.

tf_cnn_benchmarks builds every model through ``convnet_builder.ConvNetBuilder`` (``conv``, ``mpool``,
``apool``, ``affine``, ``batch_norm``, ``spatial_mean``, ``dropout`` — external to the reference tree,
selected by ``--model`` at benchmark-scripts/run-tf-sing-ucx-openmpi.sh:34,66).
This module offers the same vocabulary for PyTorch modules. Two things are different by design:

* each layer records the tf_cnn_benchmarks variable name it would have had (``v0/cg/conv0/conv2d/kernel``
  …) so checkpoints can be written under those names (utils/tfnames.py);
* each layer dispatches to either the hand-written sm_100a kernels (``kernels='b200'``, NHWC bf16) or
  stock PyTorch/cuDNN ops (``kernels='cudnn'``, also the CPU path).
"""
from __future__ import annotations

import math
from collections import defaultdict
from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple, Union

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import ops

IntPair = Union[int, Tuple[int, int]]


def _pair(v: IntPair) -> Tuple[int, int]:
    return (v, v) if isinstance(v, int) else (int(v[0]), int(v[1]))


@dataclass
class BuildConfig:
    kernels: str = "cudnn"          # 'b200' | 'cudnn'
    use_batch_norm: bool = False
    bn_decay: float = 0.999         # tf_cnn_benchmarks builder default; ResNet overrides to 0.9
    bn_eps: float = 1e-3            # ResNet overrides to 1e-5
    bn_scale: bool = False          # gamma present?
    num_classes: int = 1001         # tf_cnn_benchmarks ImageNet has a background class


ACT_MODE = {None: 0, "linear": 0, "relu": 1, "relu6": 2}     # the ``relu`` argument of the ops layer


def same_padding(size: int, k: int, s: int) -> Tuple[int, int]:
    """TensorFlow 'SAME' padding (lo, hi) for one spatial dim."""
    out = -(-size // s)
    total = max((out - 1) * s + k - size, 0)
    return total // 2, total - total // 2


class Conv(nn.Module):
    """conv (+bias) (+batch-norm) (+ReLU), the unit ``ConvNetBuilder.conv`` emits."""

    def __init__(self, cfg: BuildConfig, cin: int, cout: int, k: IntPair, stride: IntPair, mode: str,
                 in_hw: Tuple[int, int], use_bn: bool, activation: Optional[str], bias: Optional[float],
                 stddev: Optional[float], scope: str, bn_scope: str, groups: int = 1):
        super().__init__()
        self.cfg = cfg
        self.cin, self.cout = cin, cout
        self.groups = groups                      # cin == cout == groups: depthwise (MobileNet)
        assert cin % groups == 0 and cout % groups == 0
        self.k, self.stride = _pair(k), _pair(stride)
        self.activation = activation
        self.scope = scope
        self.in_hw = tuple(in_hw)
        kh, kw = self.k
        sh, sw = self.stride
        if mode == "SAME":
            ph, pw = same_padding(in_hw[0], kh, sh), same_padding(in_hw[1], kw, sw)
        elif mode == "VALID":
            ph, pw = (0, 0), (0, 0)
        elif mode == "SAME_RESNET":
            # tf_cnn_benchmarks pads explicitly by (k-1)//2 then runs VALID for strided ResNet convs.
            ph = ((kh - 1) // 2, kh - 1 - (kh - 1) // 2) if sh > 1 else same_padding(in_hw[0], kh, sh)
            pw = ((kw - 1) // 2, kw - 1 - (kw - 1) // 2) if sw > 1 else same_padding(in_hw[1], kw, sw)
        else:
            raise ValueError(mode)
        self.pad = (ph, pw)                       # ((top,bottom),(left,right))
        self.out_hw = ((in_hw[0] + ph[0] + ph[1] - kh) // sh + 1, (in_hw[1] + pw[0] + pw[1] - kw) // sw + 1)
        w = torch.empty(cout, cin // groups, kh, kw)
        if stddev is not None:
            nn.init.trunc_normal_(w, std=stddev, a=-2 * stddev, b=2 * stddev)
        else:
            # variance_scaling_initializer() default: fan_in, factor 2, truncated normal
            std = math.sqrt(2.0 / (cin // groups * kh * kw))
            nn.init.trunc_normal_(w, std=std, a=-2 * std, b=2 * std)
        self.weight = nn.Parameter(w.contiguous(memory_format=torch.channels_last))
        if groups > 1:
            # tf.nn.depthwise_conv2d keeps its filter as [kh, kw, in_channels, channel_multiplier]
            self.weight.tf_name = f"{scope}/depthwise_conv2d/depthwise_kernel"
            self.weight.tf_layout = "HWC1"
        else:
            self.weight.tf_name = f"{scope}/conv2d/kernel"
            self.weight.tf_layout = "HWIO"
        self.weight.decay = True
        if use_bn:
            self.bias = None
            self.bn = BatchNorm(cfg, cout, f"{scope}/{bn_scope}")
        else:
            self.bn = None
            if bias is not None:
                self.bias = nn.Parameter(torch.full((cout,), float(bias)))
                self.bias.tf_name = f"{scope}/biases"
                self.bias.tf_layout = "VEC"
                self.bias.decay = True   # tf_cnn_benchmarks L2-regularises every trainable variable
            else:
                self.bias = None

    def forward(self, x, residual=None):
        return ops.conv_block(x, self, relu=ACT_MODE[self.activation], residual=residual)

    def extra_repr(self):
        g = f" groups={self.groups}" if self.groups > 1 else ""
        return f"{self.cin}->{self.cout} k={self.k} s={self.stride} pad={self.pad}{g} act={self.activation} [{self.scope}]"


class BatchNorm(nn.Module):
    def __init__(self, cfg: BuildConfig, c: int, scope: str):
        super().__init__()
        self.cfg = cfg
        self.c = c
        self.eps = cfg.bn_eps
        self.decay = cfg.bn_decay
        self.scope = scope
        self.has_scale = cfg.bn_scale
        # gamma is always materialised (ones) so the kernels have one code path; it is only trainable
        # and only checkpointed when the model asks for scale=True, like the TF layer.
        self.gamma = nn.Parameter(torch.ones(c), requires_grad=cfg.bn_scale)
        self.beta = nn.Parameter(torch.zeros(c))
        self.register_buffer("moving_mean", torch.zeros(c))
        self.register_buffer("moving_variance", torch.ones(c))
        for p, n in ((self.gamma, "gamma"), (self.beta, "beta")):
            p.tf_name = f"{scope}/{n}"
            p.tf_layout = "VEC"
            p.decay = True
        self.gamma.tf_skip = not cfg.bn_scale
        self.tf_buffers = {"moving_mean": f"{scope}/moving_mean", "moving_variance": f"{scope}/moving_variance"}


class BNAct(nn.Module):
    """Standalone batch-norm (+ReLU): ``ConvNetBuilder.batch_norm`` — the pre-activation of ResNet v2 / DenseNet."""

    def __init__(self, cfg: BuildConfig, c: int, scope: str, activation: Optional[str]):
        super().__init__()
        self.cfg = cfg
        self.bn = BatchNorm(cfg, c, scope)
        self.activation = activation

    def forward(self, x):
        return ops.batch_norm(x, self.bn, relu=ACT_MODE[self.activation])


class Affine(nn.Module):
    """Fully connected layer (``ConvNetBuilder.affine``). Weights are stored (out, in); TF layout is (in, out)."""

    def __init__(self, cfg: BuildConfig, cin: int, cout: int, activation: Optional[str], bias: float,
                 stddev: Optional[float], scope: str):
        super().__init__()
        self.cfg = cfg
        self.cin, self.cout = cin, cout
        self.activation = activation
        self.scope = scope
        init_factor = 2.0 if activation == "relu" else 1.0
        std = stddev if stddev is not None else math.sqrt(init_factor / cin)
        w = torch.empty(cout, cin)
        nn.init.trunc_normal_(w, std=std, a=-2 * std, b=2 * std)
        self.weight = nn.Parameter(w)
        self.bias = nn.Parameter(torch.full((cout,), float(bias)))
        self.weight.tf_name = f"{scope}/weights"
        self.weight.tf_layout = "IO"
        self.weight.decay = True
        self.bias.tf_name = f"{scope}/biases"
        self.bias.tf_layout = "VEC"
        self.bias.decay = True

    def forward(self, x):
        return ops.affine(x, self, relu=self.activation == "relu")


class Pool(nn.Module):
    def __init__(self, cfg: BuildConfig, kind: str, k: IntPair, stride: IntPair, mode: str, in_hw):
        super().__init__()
        self.cfg = cfg
        self.kind = kind
        self.k, self.stride = _pair(k), _pair(stride)
        if mode == "SAME":
            ph, pw = same_padding(in_hw[0], self.k[0], self.stride[0]), same_padding(in_hw[1], self.k[1], self.stride[1])
        else:
            ph, pw = (0, 0), (0, 0)
        self.pad = (ph, pw)
        self.out_hw = ((in_hw[0] + sum(ph) - self.k[0]) // self.stride[0] + 1,
                       (in_hw[1] + sum(pw) - self.k[1]) // self.stride[1] + 1)

    def forward(self, x):
        return ops.pool(x, self)

    def extra_repr(self):
        return f"{self.kind} k={self.k} s={self.stride} pad={self.pad}"


class SpatialMean(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.cfg = cfg

    def forward(self, x):
        return ops.spatial_mean(x, self.cfg)


class Reshape(nn.Module):
    """Flatten in NHWC element order (what TF's reshape of an NHWC tensor yields)."""

    def __init__(self, cfg):
        super().__init__()
        self.cfg = cfg

    def forward(self, x):
        if x.dim() == 4:
            x = x.permute(0, 2, 3, 1)
        return x.reshape(x.shape[0], -1)


class Dropout(nn.Module):
    def __init__(self, keep_prob: float):
        super().__init__()
        self.p = 1.0 - keep_prob

    def forward(self, x):
        return F.dropout(x, self.p, self.training)


class ConvNetBuilder:
    """Imperative builder: tracks the running activation shape and tf_cnn_benchmarks name counters."""

    def __init__(self, cfg: BuildConfig, in_channels: int, in_hw: Tuple[int, int], prefix: str = "v0/cg"):
        self.cfg = cfg
        self.c = in_channels
        self.hw = tuple(in_hw)
        self.prefix = prefix
        self.counts = defaultdict(int)
        self.scope_stack: List[str] = []
        self.flat = False
        self.use_batch_norm = cfg.use_batch_norm

    # -- naming
    def _scope(self, key: str) -> str:
        name = f"{key}{self.counts[key]}"
        self.counts[key] += 1
        return "/".join([self.prefix] + self.scope_stack + [name])

    def push_scope(self, key: str):
        name = f"{key}{self.counts[key]}"
        self.counts[key] += 1
        self.scope_stack.append(name)

    def pop_scope(self):
        self.scope_stack.pop()

    # -- layers
    def conv(self, cout: int, k: IntPair, stride: IntPair = 1, mode: str = "SAME", cin: Optional[int] = None,
             in_hw=None, use_batch_norm: Optional[bool] = None, activation: Optional[str] = "relu",
             bias: Optional[float] = 0.0, stddev: Optional[float] = None, advance: bool = True,
             groups: int = 1) -> Conv:
        cin = self.c if cin is None else cin
        in_hw = self.hw if in_hw is None else in_hw
        use_bn = self.use_batch_norm if use_batch_norm is None else use_batch_norm
        scope = self._scope("conv")
        bn_scope = f"batchnorm{self.counts['batchnorm']}"
        if use_bn:
            self.counts["batchnorm"] += 1
        m = Conv(self.cfg, cin, cout, k, stride, mode, in_hw, use_bn, activation, bias, stddev, scope, bn_scope, groups)
        if advance:
            self.c, self.hw = cout, m.out_hw
        return m

    def batch_norm(self, activation: Optional[str] = "relu", c: Optional[int] = None) -> BNAct:
        scope = "/".join([self.prefix] + self.scope_stack + [f"batchnorm{self.counts['batchnorm']}"])
        self.counts["batchnorm"] += 1
        return BNAct(self.cfg, self.c if c is None else c, scope, activation)

    def mpool(self, k: IntPair, stride: IntPair = 2, mode: str = "VALID", in_hw=None, advance=True) -> Pool:
        self.counts["mpool"] += 1
        m = Pool(self.cfg, "max", k, stride, mode, self.hw if in_hw is None else in_hw)
        if advance:
            self.hw = m.out_hw
        return m

    def apool(self, k: IntPair, stride: IntPair = 2, mode: str = "VALID", in_hw=None, advance=True) -> Pool:
        self.counts["apool"] += 1
        m = Pool(self.cfg, "avg", k, stride, mode, self.hw if in_hw is None else in_hw)
        if advance:
            self.hw = m.out_hw
        return m

    def spatial_mean(self) -> SpatialMean:
        self.counts["spatial_mean"] += 1
        self.hw = (1, 1)
        return SpatialMean(self.cfg)

    def reshape_flat(self) -> Reshape:
        self.c = self.c * self.hw[0] * self.hw[1]
        self.hw = (1, 1)
        self.flat = True
        return Reshape(self.cfg)

    def affine(self, cout: int, activation: Optional[str] = "relu", bias: float = 0.0,
               stddev: Optional[float] = None) -> Affine:
        scope = self._scope("affine")
        m = Affine(self.cfg, self.c, cout, activation, bias, stddev, scope)
        self.c = cout
        return m

    def dropout(self, keep_prob: float = 0.5) -> Dropout:
        self.counts["dropout"] += 1
        return Dropout(keep_prob)

    def inception_module(self, cols: Sequence[Sequence[tuple]]) -> "InceptionModule":
        """tf_cnn_benchmarks ``inception_module``: ``cols`` is a list of branches, each a list of
        ``('conv', cout, kh, kw[, sh, sw, mode])``, ``('mpool'|'apool', kh, kw, sh, sw[, mode])`` or
        ``('share',)`` entries. ``('share',)`` at depth i of branch b reuses the *output* of branch b-1's
        layer i (computed once), as in TF."""
        self.push_scope("incept")
        in_c, in_hw = self.c, self.hw
        built: List[List[Optional[nn.Module]]] = []
        shapes: List[List[Tuple[int, Tuple[int, int]]]] = []
        out_c, out_hw = 0, None
        for b, col in enumerate(cols):
            layers: List[Optional[nn.Module]] = []
            shp = []
            c, hw = in_c, in_hw
            for i, spec in enumerate(col):
                kind = spec[0]
                if kind == "share":
                    c, hw = shapes[b - 1][i]
                    layers.append(None)
                elif kind == "conv":
                    cout, kh, kw = spec[1], spec[2], spec[3]
                    sh, sw = (spec[4], spec[5]) if len(spec) > 5 else (1, 1)
                    mode = spec[6] if len(spec) > 6 else "SAME"
                    m = self.conv(cout, (kh, kw), (sh, sw), mode, cin=c, in_hw=hw, advance=False)
                    c, hw = cout, m.out_hw
                    layers.append(m)
                elif kind in ("mpool", "apool"):
                    kh, kw, sh, sw = spec[1], spec[2], spec[3], spec[4]
                    mode = spec[5] if len(spec) > 5 else "SAME"
                    fn = self.mpool if kind == "mpool" else self.apool
                    m = fn((kh, kw), (sh, sw), mode, in_hw=hw, advance=False)
                    hw = m.out_hw
                    layers.append(m)
                else:
                    raise ValueError(spec)
                shp.append((c, hw))
            built.append(layers)
            shapes.append(shp)
            out_c += c
            out_hw = hw
        self.c, self.hw = out_c, out_hw
        self.pop_scope()
        return InceptionModule(built)


class InceptionModule(nn.Module):
    def __init__(self, cols):
        super().__init__()
        self.share = [[m is None for m in col] for col in cols]
        self.cols = nn.ModuleList(nn.ModuleList([m if m is not None else nn.Identity() for m in col])
                                  for col in cols)

    def forward(self, x):
        outs = []
        prev: List[torch.Tensor] = []
        for b, col in enumerate(self.cols):
            cur = []
            t = x
            for i, m in enumerate(col):
                t = prev[i] if self.share[b][i] else m(t)
                cur.append(t)
            outs.append(t)
            prev = cur
        return ops.concat_channels(outs)


class Model(nn.Module):
    """Base class: ``body`` is a list of modules applied in order; ``logits`` is the final affine."""

    name = "model"
    image_size = 224
    default_batch_size = 64
    default_learning_rate = 0.005

    def __init__(self, cfg: BuildConfig):
        super().__init__()
        self.cfg = cfg

    def get_learning_rate(self, global_batch: int) -> float:
        return self.default_learning_rate

    def features(self, x):
        raise NotImplementedError

    def forward(self, x):
        return self.head(self.features(x))
