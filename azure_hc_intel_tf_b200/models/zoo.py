"""VGG, Inception-v3, GoogLeNet, AlexNet, trivial (and, in zoo2.py, ResNet v2, the CIFAR-10 ResNets / DenseNets and
Inception-v4) — the rest of the tf_cnn_benchmarks zoo the harness can
name with ``--model`` (the reference hard-codes resnet50 at run-tf-sing-ucx-openmpi.sh:34 and tells the
user to edit the script for others; BASELINE.json configs add VGG-16 and Inception-v3).
"""
from __future__ import annotations

import torch.nn as nn

from .builder import BuildConfig, ConvNetBuilder, Model


class Sequential(Model):
    """A model that is a straight chain of builder layers followed by the classifier head."""

    def _finish(self, b: ConvNetBuilder, layers):
        self.body = nn.ModuleList(layers)
        self.head = b.affine(self.cfg.num_classes, activation=None)

    def features(self, x):
        for m in self.body:
            x = m(x)
        return x


class VGG(Sequential):
    image_size = 224
    default_batch_size = 64
    default_learning_rate = 0.005

    def __init__(self, cfg: BuildConfig, num_conv_layers, name):
        super().__init__(cfg)
        self.name = name
        b = ConvNetBuilder(cfg, 3, (224, 224))
        layers = []
        for n, width in zip(num_conv_layers, (64, 128, 256, 512, 512)):
            for _ in range(n):
                layers.append(b.conv(width, 3))
            layers.append(b.mpool(2, 2))
        layers.append(b.reshape_flat())
        layers += [b.affine(4096), b.dropout(), b.affine(4096), b.dropout()]
        self._finish(b, layers)


class AlexNet(Sequential):
    image_size = 227
    default_batch_size = 512
    default_learning_rate = 0.005

    def __init__(self, cfg):
        super().__init__(cfg)
        self.name = "alexnet"
        b = ConvNetBuilder(cfg, 3, (227, 227))
        layers = [b.conv(64, 11, 4, "VALID"), b.mpool(3, 2), b.conv(192, 5), b.mpool(3, 2),
                  b.conv(384, 3), b.conv(384, 3), b.conv(256, 3), b.mpool(3, 2), b.reshape_flat(),
                  b.affine(4096), b.dropout(), b.affine(4096), b.dropout()]
        self._finish(b, layers)


class LeNet(Sequential):
    """tf_cnn_benchmarks lenet5 (28x28 input)."""
    image_size = 28
    default_batch_size = 32

    def __init__(self, cfg):
        super().__init__(cfg)
        self.name = "lenet"
        b = ConvNetBuilder(cfg, 3, (28, 28))
        layers = [b.conv(32, 5), b.mpool(2, 2), b.conv(64, 5), b.mpool(2, 2), b.reshape_flat(), b.affine(512)]
        self._finish(b, layers)


class Overfeat(Sequential):
    image_size = 231
    default_batch_size = 32

    def __init__(self, cfg):
        super().__init__(cfg)
        self.name = "overfeat"
        b = ConvNetBuilder(cfg, 3, (231, 231))
        layers = [b.conv(96, 11, 4, "VALID"), b.mpool(2, 2), b.conv(256, 5, 1, "VALID"), b.mpool(2, 2),
                  b.conv(512, 3), b.conv(1024, 3), b.conv(1024, 3), b.mpool(2, 2), b.reshape_flat(),
                  b.affine(3072), b.dropout(), b.affine(4096), b.dropout()]
        self._finish(b, layers)


class Trivial(Sequential):
    image_size = 227
    default_batch_size = 32

    def __init__(self, cfg):
        super().__init__(cfg)
        self.name = "trivial"
        b = ConvNetBuilder(cfg, 3, (227, 227))
        layers = [b.reshape_flat(), b.affine(1), b.affine(4096)]
        self._finish(b, layers)


class GoogLeNet(Sequential):
    image_size = 224
    default_batch_size = 32
    default_learning_rate = 0.005

    def __init__(self, cfg):
        super().__init__(cfg)
        self.name = "googlenet"
        b = ConvNetBuilder(cfg, 3, (224, 224))

        def v1(k, l, m, n, p, q):
            return b.inception_module([[("conv", k, 1, 1)], [("conv", l, 1, 1), ("conv", m, 3, 3)],
                                       [("conv", n, 1, 1), ("conv", p, 5, 5)],
                                       [("mpool", 3, 3, 1, 1, "SAME"), ("conv", q, 1, 1)]])

        layers = [b.conv(64, 7, 2), b.mpool(3, 2, "SAME"), b.conv(64, 1), b.conv(192, 3), b.mpool(3, 2, "SAME"),
                  v1(64, 96, 128, 16, 32, 32), v1(128, 128, 192, 32, 96, 64), b.mpool(3, 2, "SAME"),
                  v1(192, 96, 208, 16, 48, 64), v1(160, 112, 224, 24, 64, 64), v1(128, 128, 256, 24, 64, 64),
                  v1(112, 144, 288, 32, 64, 64), v1(256, 160, 320, 32, 128, 128), b.mpool(3, 2, "SAME"),
                  v1(256, 160, 320, 32, 128, 128), v1(384, 192, 384, 48, 128, 128),
                  b.apool(7, 1, "VALID"), b.reshape_flat()]
        self._finish(b, layers)


class InceptionV3(Sequential):
    image_size = 299
    default_batch_size = 32
    default_learning_rate = 0.005

    def __init__(self, cfg):
        cfg.use_batch_norm = True   # builder defaults: decay 0.999, eps 1e-3, no gamma
        super().__init__(cfg)
        self.name = "inception3"
        b = ConvNetBuilder(cfg, 3, (299, 299))
        im = b.inception_module

        def a(n):
            return im([[("conv", 64, 1, 1)], [("conv", 48, 1, 1), ("conv", 64, 5, 5)],
                       [("conv", 64, 1, 1), ("conv", 96, 3, 3), ("conv", 96, 3, 3)],
                       [("apool", 3, 3, 1, 1, "SAME"), ("conv", n, 1, 1)]])

        def bmod():
            return im([[("conv", 384, 3, 3, 2, 2, "VALID")],
                       [("conv", 64, 1, 1), ("conv", 96, 3, 3), ("conv", 96, 3, 3, 2, 2, "VALID")],
                       [("mpool", 3, 3, 2, 2, "VALID")]])

        def c(n):
            return im([[("conv", 192, 1, 1)],
                       [("conv", n, 1, 1), ("conv", n, 1, 7), ("conv", 192, 7, 1)],
                       [("conv", n, 1, 1), ("conv", n, 7, 1), ("conv", n, 1, 7), ("conv", n, 7, 1),
                        ("conv", 192, 1, 7)],
                       [("apool", 3, 3, 1, 1, "SAME"), ("conv", 192, 1, 1)]])

        def d():
            return im([[("conv", 192, 1, 1), ("conv", 320, 3, 3, 2, 2, "VALID")],
                       [("conv", 192, 1, 1), ("conv", 192, 1, 7), ("conv", 192, 7, 1),
                        ("conv", 192, 3, 3, 2, 2, "VALID")],
                       [("mpool", 3, 3, 2, 2, "VALID")]])

        def e(pooltype):
            return im([[("conv", 320, 1, 1)], [("conv", 384, 1, 1), ("conv", 384, 1, 3)],
                       [("share",), ("conv", 384, 3, 1)],
                       [("conv", 448, 1, 1), ("conv", 384, 3, 3), ("conv", 384, 1, 3)],
                       [("share",), ("share",), ("conv", 384, 3, 1)],
                       [("mpool" if pooltype == "max" else "apool", 3, 3, 1, 1, "SAME"), ("conv", 192, 1, 1)]])

        layers = [b.conv(32, 3, 2, "VALID"), b.conv(32, 3, 1, "VALID"), b.conv(64, 3, 1, "SAME"),
                  b.mpool(3, 2, "VALID"), b.conv(80, 1, 1, "VALID"), b.conv(192, 3, 1, "VALID"),
                  b.mpool(3, 2, "VALID"),
                  a(32), a(64), a(64), bmod(), c(128), c(160), c(160), c(192), d(), e("avg"), e("max"),
                  b.apool(8, 1, "VALID"), b.reshape_flat()]
        self._finish(b, layers)
