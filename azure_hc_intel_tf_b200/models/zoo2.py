"""More of the tf_cnn_benchmarks zoo: pre-activation ResNets (v2), the CIFAR-10 ResNets and DenseNets, Inception-v4.

The reference hard-codes ``MODEL=resnet50`` (benchmark-scripts/run-tf-sing-ucx-openmpi.sh:34) and tells the user to
edit the script for anything else; ``--model`` of tf_cnn_benchmarks accepts these names (model zoo
row). They are built from the same ConvNetBuilder pieces as the rest of the zoo, so variable names follow the
tf_cnn_benchmarks scopes (``v0/cg/...``) and every layer dispatches through ``ops`` (own sm_100a kernels where the
channel counts allow, the stock ops otherwise - DenseNet's growth rate of 12 gives channel counts that are not a
multiple of the 8-channel vectors the kernels move).
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import ops
from .builder import BuildConfig, ConvNetBuilder, Model


# ------------------------------------------------------------------------------------ ResNet v2 (ImageNet)

class BottleneckV2(nn.Module):
    """tf_cnn_benchmarks ``bottleneck_block_v2``: BN+ReLU pre-activation, three convs, no activation after the sum."""

    def __init__(self, b: ConvNetBuilder, depth: int, depth_bottleneck: int, stride: int):
        super().__init__()
        self.cfg = b.cfg
        in_c, in_hw = b.c, b.hw
        b.push_scope("resnet_v2")
        self.preact = b.batch_norm("relu")
        if depth == in_c:
            self.shortcut = None if stride == 1 else b.apool(1, stride, "VALID", in_hw=in_hw, advance=False)
            self.shortcut_from_preact = False
        else:
            self.shortcut = b.conv(depth, 1, stride, "SAME", cin=in_c, in_hw=in_hw, use_batch_norm=False,
                                   activation=None, bias=None, advance=False)
            self.shortcut_from_preact = True
        self.conv1 = b.conv(depth_bottleneck, 1, stride, "SAME", cin=in_c, in_hw=in_hw, use_batch_norm=True, bias=None)
        self.conv2 = b.conv(depth_bottleneck, 3, 1, "SAME", use_batch_norm=True, bias=None)
        self.conv3 = b.conv(depth, 1, 1, "SAME", use_batch_norm=False, activation=None, bias=None)
        b.pop_scope()

    def forward(self, x):
        pre = self.preact(x)
        if self.shortcut is None:
            sc = x
        else:
            sc = self.shortcut(pre if self.shortcut_from_preact else x)
        return ops.add(sc, self.conv3(self.conv2(self.conv1(pre))), self.cfg)


class ResNetV2(Model):
    image_size = 224
    default_batch_size = 64

    def __init__(self, cfg: BuildConfig, layer_counts, name: str):
        cfg.use_batch_norm = True
        cfg.bn_decay, cfg.bn_eps, cfg.bn_scale = 0.9, 1e-5, True
        super().__init__(cfg)
        self.name = name
        b = ConvNetBuilder(cfg, 3, (224, 224))
        self.conv0 = b.conv(64, 7, 2, "SAME_RESNET", use_batch_norm=True)
        self.pool0 = b.mpool(3, 2, "SAME")
        blocks = []
        for stage, (depth, bott) in enumerate(((256, 64), (512, 128), (1024, 256), (2048, 512))):
            for i in range(layer_counts[stage]):
                blocks.append(BottleneckV2(b, depth, bott, 2 if (i == 0 and stage > 0) else 1))
        self.blocks = nn.ModuleList(blocks)
        self.postact = b.batch_norm("relu")
        self.gap = b.spatial_mean()
        self.head = b.affine(cfg.num_classes, activation=None)

    def features(self, x):
        x = self.pool0(self.conv0(x))
        for blk in self.blocks:
            x = blk(x)
        return self.gap(self.postact(x))

    def get_learning_rate(self, global_batch: int) -> float:
        return 0.128 * global_batch / 256.0


# ------------------------------------------------------------------------------------ CIFAR-10 ResNets

class ResidualBlockCifar(nn.Module):
    """tf_cnn_benchmarks ``residual_block``: two 3x3 convs; when the width changes the shortcut is a strided average
    pool followed by zero padding of the channels (option A of the ResNet paper), so it has no parameters."""

    def __init__(self, b: ConvNetBuilder, depth: int, stride: int, v2: bool):
        super().__init__()
        self.cfg = b.cfg
        self.v2 = v2
        in_c, in_hw = b.c, b.hw
        self.in_c, self.depth = in_c, depth
        b.push_scope("resnet_v2" if v2 else "resnet_v1")
        self.pool = b.apool(stride, stride, "VALID", in_hw=in_hw, advance=False) if (in_c != depth and stride > 1) else None
        if v2:
            self.preact = b.batch_norm("relu")
            self.conv1 = b.conv(depth, 3, stride, "SAME_RESNET", use_batch_norm=True, bias=None)
            self.conv2 = b.conv(depth, 3, 1, "SAME_RESNET", use_batch_norm=False, activation=None, bias=None)
        else:
            self.preact = None
            self.conv1 = b.conv(depth, 3, stride, "SAME_RESNET", use_batch_norm=True, bias=None)
            self.conv2 = b.conv(depth, 3, 1, "SAME_RESNET", use_batch_norm=True, activation=None, bias=None)
        b.pop_scope()

    def forward(self, x):
        sc = x
        if self.in_c != self.depth:
            if self.pool is not None:
                sc = self.pool(sc)
            lo = (self.depth - self.in_c) // 2
            sc = F.pad(sc, (0, 0, 0, 0, lo, self.depth - self.in_c - lo))
            sc = sc.contiguous(memory_format=torch.channels_last)
        h = self.preact(x) if self.v2 else x
        y = self.conv2(self.conv1(h))
        return ops.add(sc, y, self.cfg) if self.v2 else ops.add_relu(sc, y, self.cfg)


class ResNetCifar(Model):
    image_size = 32
    default_batch_size = 128
    default_learning_rate = 0.1
    num_classes_default = 11          # tf_cnn_benchmarks counts an unused background class for every dataset

    def __init__(self, cfg: BuildConfig, depth: int, v2: bool, name: str):
        assert (depth - 2) % 6 == 0
        n = (depth - 2) // 6
        cfg.use_batch_norm = True
        cfg.bn_decay, cfg.bn_eps, cfg.bn_scale = 0.9, 1e-5, True
        super().__init__(cfg)
        self.name = name
        b = ConvNetBuilder(cfg, 3, (32, 32))
        self.conv0 = b.conv(16, 3, 1, "SAME_RESNET", use_batch_norm=not v2, activation=None if v2 else "relu", bias=None)
        blocks = []
        for stage, width in enumerate((16, 32, 64)):
            for i in range(n):
                blocks.append(ResidualBlockCifar(b, width, 2 if (i == 0 and stage > 0) else 1, v2))
        self.blocks = nn.ModuleList(blocks)
        self.postact = b.batch_norm("relu") if v2 else None
        self.gap = b.spatial_mean()
        self.head = b.affine(cfg.num_classes, activation=None)

    def features(self, x):
        x = self.conv0(x)
        for blk in self.blocks:
            x = blk(x)
        if self.postact is not None:
            x = self.postact(x)
        return self.gap(x)


# ------------------------------------------------------------------------------------ DenseNet (CIFAR-10)

class DenseLayer(nn.Module):
    def __init__(self, b: ConvNetBuilder, growth: int):
        super().__init__()
        self.bn = b.batch_norm("relu")
        self.conv = b.conv(growth, 3, 1, "SAME", use_batch_norm=False, activation=None, bias=None,
                           stddev=math.sqrt(2.0 / (9 * growth)), advance=False)
        b.c += growth

    def forward(self, x):
        return ops.concat_channels([x, self.conv(self.bn(x))])


class Transition(nn.Module):
    def __init__(self, b: ConvNetBuilder):
        super().__init__()
        self.bn = b.batch_norm("relu")
        self.conv = b.conv(b.c, 1, 1, "SAME", use_batch_norm=False, activation=None, bias=None,
                           stddev=math.sqrt(2.0 / b.c))
        self.pool = b.apool(2, 2)

    def forward(self, x):
        return self.pool(self.conv(self.bn(x)))


class DenseNetCifar(Model):
    image_size = 32
    default_batch_size = 64
    default_learning_rate = 0.1
    num_classes_default = 11

    def __init__(self, cfg: BuildConfig, layer_counts, growth: int, name: str):
        cfg.use_batch_norm = False          # batch-norms are explicit pre-activations here
        cfg.bn_decay, cfg.bn_eps, cfg.bn_scale = 0.9, 1e-5, True
        super().__init__(cfg)
        self.name = name
        b = ConvNetBuilder(cfg, 3, (32, 32))
        layers = [b.conv(16, 3, 1, "SAME", use_batch_norm=False, activation=None, bias=None)]
        for stage, count in enumerate(layer_counts):
            layers += [DenseLayer(b, growth) for _ in range(count)]
            if stage < len(layer_counts) - 1:
                layers.append(Transition(b))
        layers.append(b.batch_norm("relu"))
        self.body = nn.ModuleList(layers)
        self.gap = b.spatial_mean()         # tf_cnn_benchmarks: apool over the remaining 8x8 map, then reshape
        self.head = b.affine(cfg.num_classes, activation=None)

    def features(self, x):
        for m in self.body:
            x = m(x)
        return self.gap(x)


# ------------------------------------------------------------------------------------ Inception-v4

class InceptionV4(Model):
    image_size = 299
    default_batch_size = 32
    default_learning_rate = 0.005

    def __init__(self, cfg: BuildConfig):
        cfg.use_batch_norm = True
        super().__init__(cfg)
        self.name = "inception4"
        b = ConvNetBuilder(cfg, 3, (299, 299))
        im = b.inception_module

        def stem():
            return [b.conv(32, 3, 2, "VALID"), b.conv(32, 3, 1, "VALID"), b.conv(64, 3),
                    im([[("mpool", 3, 3, 2, 2, "VALID")], [("conv", 96, 3, 3, 2, 2, "VALID")]]),
                    im([[("conv", 64, 1, 1), ("conv", 96, 3, 3, 1, 1, "VALID")],
                        [("conv", 64, 1, 1), ("conv", 64, 7, 1), ("conv", 64, 1, 7), ("conv", 96, 3, 3, 1, 1, "VALID")]]),
                    im([[("conv", 192, 3, 3, 2, 2, "VALID")], [("mpool", 3, 3, 2, 2, "VALID")]])]

        def a():
            return im([[("apool", 3, 3, 1, 1, "SAME"), ("conv", 96, 1, 1)], [("conv", 96, 1, 1)],
                       [("conv", 64, 1, 1), ("conv", 96, 3, 3)],
                       [("conv", 64, 1, 1), ("conv", 96, 3, 3), ("conv", 96, 3, 3)]])

        def bmod():
            return im([[("apool", 3, 3, 1, 1, "SAME"), ("conv", 128, 1, 1)], [("conv", 384, 1, 1)],
                       [("conv", 192, 1, 1), ("conv", 224, 1, 7), ("conv", 256, 7, 1)],
                       [("conv", 192, 1, 1), ("conv", 192, 1, 7), ("conv", 224, 7, 1), ("conv", 224, 1, 7),
                        ("conv", 256, 7, 1)]])

        def c():
            return im([[("apool", 3, 3, 1, 1, "SAME"), ("conv", 256, 1, 1)], [("conv", 256, 1, 1)],
                       [("conv", 384, 1, 1), ("conv", 256, 1, 3)],
                       [("share",), ("conv", 256, 3, 1)],
                       [("conv", 384, 1, 1), ("conv", 448, 1, 3), ("conv", 512, 3, 1), ("conv", 256, 3, 1)],
                       [("share",), ("share",), ("share",), ("conv", 256, 1, 3)]])

        def reduction_a(k, l, m, n):
            return im([[("mpool", 3, 3, 2, 2, "VALID")], [("conv", n, 3, 3, 2, 2, "VALID")],
                       [("conv", k, 1, 1), ("conv", l, 3, 3), ("conv", m, 3, 3, 2, 2, "VALID")]])

        def reduction_b():
            return im([[("mpool", 3, 3, 2, 2, "VALID")],
                       [("conv", 192, 1, 1), ("conv", 192, 3, 3, 2, 2, "VALID")],
                       [("conv", 256, 1, 1), ("conv", 256, 1, 7), ("conv", 320, 7, 1), ("conv", 320, 3, 3, 2, 2, "VALID")]])

        layers = stem()
        layers += [a() for _ in range(4)]
        layers.append(reduction_a(192, 224, 256, 384))
        layers += [bmod() for _ in range(7)]
        layers.append(reduction_b())
        layers += [c() for _ in range(3)]
        layers.append(b.spatial_mean())
        layers.append(b.dropout(0.8))
        self.body = nn.ModuleList(layers)
        self.head = b.affine(cfg.num_classes, activation=None)

    def features(self, x):
        for m in self.body:
            x = m(x)
        return x
