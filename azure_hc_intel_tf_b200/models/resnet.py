"""ResNet v1 / v1.5 (50, 101, 152) in tf_cnn_benchmarks form.

``--model=resnet50`` is the reference's only configured model
(benchmark-scripts/run-tf-sing-ucx-openmpi.sh:34). tf_cnn_benchmarks' ``resnet50`` is v1: the
down-sampling stride sits in the first 1×1 convolution of a stage's first block; ``resnet50_v1.5``
moves it to the 3×3. BatchNorm uses decay 0.9, epsilon 1e-5, scale=True; the head is a 1001-way
affine layer.
"""
from __future__ import annotations

import torch.nn as nn

from .. import ops
from .builder import BuildConfig, ConvNetBuilder, Model


class Bottleneck(nn.Module):
    def __init__(self, b: ConvNetBuilder, depth: int, depth_bottleneck: int, stride: int, v15: bool):
        super().__init__()
        self.cfg = b.cfg
        in_c, in_hw = b.c, b.hw
        b.push_scope("resnet_v1")
        if depth == in_c:
            if stride == 1:
                self.shortcut = None
            else:
                self.shortcut = b.apool(1, stride, "VALID", in_hw=in_hw, advance=False)
        else:
            self.shortcut = b.conv(depth, 1, stride, "SAME", cin=in_c, in_hw=in_hw, use_batch_norm=True,
                                   activation=None, bias=None, advance=False)
        s1, s3 = (1, stride) if v15 else (stride, 1)
        self.conv1 = b.conv(depth_bottleneck, 1, s1, "SAME", cin=in_c, in_hw=in_hw, use_batch_norm=True, bias=None)
        self.conv2 = b.conv(depth_bottleneck, 3, s3, "SAME_RESNET", use_batch_norm=True, bias=None)
        self.conv3 = b.conv(depth, 1, 1, "SAME", use_batch_norm=True, activation=None, bias=None)
        b.pop_scope()

    def forward(self, x):
        return ops.bottleneck(x, self)


class BasicBlock(nn.Module):
    """Two 3x3 convolutions (ResNet-18 / -34 of the TF official models that tf_cnn_benchmarks exposes as
    ``official_resnet18`` / ``official_resnet34``): relu(bn2(conv2(relu(bn1(conv1(x))))) + shortcut(x)), projection
    shortcut (1x1, stride 2, BN) where the shape changes. The sum and the output ReLU ride in the second BN's apply."""

    def __init__(self, b: ConvNetBuilder, depth: int, stride: int):
        super().__init__()
        self.cfg = b.cfg
        in_c, in_hw = b.c, b.hw
        b.push_scope("resnet_v1")
        if depth == in_c and stride == 1:
            self.shortcut = None
        else:
            self.shortcut = b.conv(depth, 1, stride, "SAME", cin=in_c, in_hw=in_hw, use_batch_norm=True,
                                   activation=None, bias=None, advance=False)
        self.conv1 = b.conv(depth, 3, stride, "SAME_RESNET", cin=in_c, in_hw=in_hw, use_batch_norm=True, bias=None)
        self.conv2 = b.conv(depth, 3, 1, "SAME_RESNET", use_batch_norm=True, activation=None, bias=None)
        b.pop_scope()

    def forward(self, x):
        sc = x if self.shortcut is None else self.shortcut(x)
        return ops.conv_bn_add_relu(self.conv1(x), self.conv2, sc)


class ResNet(Model):
    image_size = 224
    default_batch_size = 64

    def __init__(self, cfg: BuildConfig, layer_counts, v15: bool = False, name: str = "resnet50", basic: bool = False):
        cfg.use_batch_norm = True
        cfg.bn_decay, cfg.bn_eps, cfg.bn_scale = 0.9, 1e-5, True
        super().__init__(cfg)
        self.name = name
        b = ConvNetBuilder(cfg, 3, (self.image_size, self.image_size))
        self.conv0 = b.conv(64, 7, 2, "SAME_RESNET", use_batch_norm=True)
        self.pool0 = b.mpool(3, 2, "SAME")
        blocks = []
        for stage, (depth, bott) in enumerate(((256, 64), (512, 128), (1024, 256), (2048, 512))):
            for i in range(layer_counts[stage]):
                stride = 2 if (i == 0 and stage > 0) else 1
                blocks.append(BasicBlock(b, bott, stride) if basic else Bottleneck(b, depth, bott, stride, v15))
        self.blocks = nn.ModuleList(blocks)
        self.gap = b.spatial_mean()
        self.head = b.affine(cfg.num_classes, activation=None)

    def features(self, x):
        x = self.pool0(self.conv0(x))
        for blk in self.blocks:
            x = blk(x)
        return self.gap(x)

    def get_learning_rate(self, global_batch: int) -> float:
        return 0.128 * global_batch / 256.0


def resnet50(cfg):
    return ResNet(cfg, (3, 4, 6, 3), False, "resnet50")


def resnet50_v15(cfg):
    return ResNet(cfg, (3, 4, 6, 3), True, "resnet50_v1.5")


def resnet101(cfg):
    return ResNet(cfg, (3, 4, 23, 3), False, "resnet101")


def resnet152(cfg):
    return ResNet(cfg, (3, 8, 36, 3), False, "resnet152")


def official_resnet18(cfg):
    return ResNet(cfg, (2, 2, 2, 2), False, "official_resnet18", basic=True)


def official_resnet34(cfg):
    return ResNet(cfg, (3, 4, 6, 3), False, "official_resnet34", basic=True)
