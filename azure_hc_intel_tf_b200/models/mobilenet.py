"""MobileNet-v2 — tf_cnn_benchmarks' ``--model=mobilenet`` (models/mobilenet_v2.py wrapping slim's definition).

The reference hard-codes ``MODEL=resnet50`` (benchmark-scripts/run-tf-sing-ucx-openmpi.sh:34) and passes it as
``--model`` (:66); MobileNet is one of the names that flag accepts (model zoo row). It is the one
family of the zoo built on depthwise convolutions and ReLU6, which is why the ops layer has both:

* 3x3 depthwise layers run on the vector kernels of csrc/kernels/depthwise.cu (bandwidth-bound: 9 multiply-adds per
  element, no GEMM in them), followed by the same BN kernels as everywhere else;
* the 1x1 expansion / projection layers are ordinary tensor-core implicit GEMMs (csrc/conv);
* ReLU6 is ``relu=2`` of the BN-apply kernel, its pass-through set (0 < f < 6) kept as one bit per element for the
  backward; the linear-bottleneck residual add is fused into the projection layer's BN-apply (``residual=``).

Architecture (Sandler et al. 2018, width 1.0, 224x224): 3x3/2 conv 32, then inverted-residual stages
(t, c, n, s) = (1,16,1,1) (6,24,2,2) (6,32,3,2) (6,64,4,2) (6,96,3,1) (6,160,3,2) (6,320,1,1), 1x1 conv 1280, global
average pool, dropout, 1001-way classifier. Every conv is followed by BN (decay 0.997, eps 1e-3, with scale).
"""
from __future__ import annotations

import torch.nn as nn

from .builder import BuildConfig, ConvNetBuilder, Model

STAGES = ((1, 16, 1, 1), (6, 24, 2, 2), (6, 32, 3, 2), (6, 64, 4, 2), (6, 96, 3, 1), (6, 160, 3, 2), (6, 320, 1, 1))


class InvertedResidual(nn.Module):
    """slim ``expanded_conv``: 1x1 expand (skipped when t == 1) -> 3x3 depthwise -> linear 1x1 projection, plus the
    input when the block keeps shape."""

    def __init__(self, b: ConvNetBuilder, expand: int, cout: int, stride: int):
        super().__init__()
        cin = b.c
        self.use_res = stride == 1 and cin == cout
        b.push_scope("expanded_conv")
        kw = dict(use_batch_norm=True, bias=None, stddev=0.09)
        self.expand = b.conv(cin * expand, 1, 1, "SAME", activation="relu6", **kw) if expand != 1 else None
        self.depthwise = b.conv(b.c, 3, stride, "SAME", activation="relu6", groups=b.c, **kw)
        self.project = b.conv(cout, 1, 1, "SAME", activation=None, **kw)
        b.pop_scope()

    def forward(self, x):
        h = x if self.expand is None else self.expand(x)
        return self.project(self.depthwise(h), residual=x if self.use_res else None)


class MobileNetV2(Model):
    image_size = 224
    default_batch_size = 32
    default_learning_rate = 0.045

    def __init__(self, cfg: BuildConfig, name: str = "mobilenet"):
        cfg.use_batch_norm = True
        cfg.bn_decay, cfg.bn_eps, cfg.bn_scale = 0.997, 1e-3, True
        super().__init__(cfg)
        self.name = name
        b = ConvNetBuilder(cfg, 3, (224, 224))
        kw = dict(use_batch_norm=True, bias=None, stddev=0.09, activation="relu6")
        self.conv0 = b.conv(32, 3, 2, "SAME", **kw)
        blocks = []
        for t, c, n, s in STAGES:
            for i in range(n):
                blocks.append(InvertedResidual(b, t, c, s if i == 0 else 1))
        self.blocks = nn.ModuleList(blocks)
        self.conv_last = b.conv(1280, 1, 1, "SAME", **kw)
        self.gap = b.spatial_mean()
        self.drop = b.dropout(0.8)
        self.head = b.affine(cfg.num_classes, activation=None)

    def features(self, x):
        x = self.conv0(x)
        for blk in self.blocks:
            x = blk(x)
        return self.drop(self.gap(self.conv_last(x)))

    def get_learning_rate(self, global_batch: int) -> float:
        # slim's recipe: 0.045 per 96 images, decayed 0.98 per epoch (utils/lr_schedule.py provides the schedules)
        return 0.045 * global_batch / 96.0
