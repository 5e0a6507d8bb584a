"""NASNet-A (mobile and large) — tf_cnn_benchmarks' ``--model=nasnet`` / ``nasnetlarge`` (models/nasnet_model.py wrapping
slim's definition; the flag is passed at benchmark-scripts/run-tf-sing-ucx-openmpi.sh:66, SURVEY.md §2.3 model zoo row).

Architecture (Zoph et al. 2018, NASNet-A): a 3x3/2 VALID stem, two reduction cells as the rest of the stem, then three
stages of N normal cells separated by reduction cells, ReLU, global average pool, dropout, classifier. Every cell takes
the previous cell's output ``ip`` and the one before it ``p``; ``p`` is first brought to the cell's shape (1x1 conv, or
the two-path factorized reduction when its resolution differs). The cell bodies are sums of separable convolutions
(ReLU, depthwise k x k, pointwise 1x1, BN - twice), 3x3 average / max pools and identities, concatenated.

Mobile: N = 4, 1056 penultimate filters (44 per branch in the first stage), 32 stem filters, 224x224 -
5,289,978 trainable parameters with 1000 classes (5,291,035 with tf_cnn_benchmarks' 1001), which the tests check.
Large: N = 6, 4032 penultimate filters, 96 stem filters, 331x331. The auxiliary head slim trains with is left out.

Kernel path: 3x3 depthwise layers whose channel count is a multiple of 8 run on csrc/kernels/depthwise.cu, 5x5 / 7x7
ones on its large-filter kernels, pointwise layers on the tensor-core implicit GEMM; the first stage of the mobile
model (11 / 22 / 44 channels, not multiples of the 8-channel vectors) takes the stock ops, like DenseNet's growth 12.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.nn as nn

from .. import ops
from .builder import BuildConfig, ConvNetBuilder, Model

Shape = Tuple[int, Tuple[int, int]]          # (channels, (h, w))


class Relu(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.cfg = cfg

    def forward(self, x):
        return ops.relu(x, self.cfg)


class SepConv(nn.Module):
    """slim ``_stacked_separable_conv``: (ReLU, depthwise k x k, pointwise 1x1, BN) twice; the stride sits in the first."""

    def __init__(self, b: ConvNetBuilder, inp: Shape, filters: int, k: int, stride: int):
        super().__init__()
        c, hw = inp
        plain = dict(use_batch_norm=False, activation=None, bias=None, advance=False)
        self.relu = Relu(b.cfg)
        self.dw1 = b.conv(c, k, stride, "SAME", cin=c, in_hw=hw, groups=c, **plain)
        self.pw1 = b.conv(filters, 1, 1, "SAME", cin=c, in_hw=self.dw1.out_hw, use_batch_norm=True, activation="relu",
                          bias=None, advance=False)
        self.dw2 = b.conv(filters, k, 1, "SAME", cin=filters, in_hw=self.dw1.out_hw, groups=filters, **plain)
        self.pw2 = b.conv(filters, 1, 1, "SAME", cin=filters, in_hw=self.dw1.out_hw, use_batch_norm=True, activation=None,
                          bias=None, advance=False)
        self.out: Shape = (filters, self.dw1.out_hw)

    def forward(self, x):
        return self.pw2(self.dw2(self.pw1(self.dw1(self.relu(x)))))


class Adjust(nn.Module):
    """Bring ``p`` (the cell input before last) to ``filters`` channels at the resolution of ``ip``."""

    def __init__(self, b: ConvNetBuilder, p: Optional[Shape], ip: Shape, filters: int):
        super().__init__()
        self.cfg = b.cfg
        self.mode = "same"
        self.out: Shape = ip
        if p is None:            # first cell: p is the cell input itself, left as it is (slim _reduce_prev_layer)
            return
        pc, phw = p
        if phw != ip[1]:
            # factorized reduction: two stride-2 1x1 average pools, the second shifted by one pixel, half the filters each
            self.mode = "reduce"
            self.relu = Relu(b.cfg)
            self.pool1 = b.apool(1, 2, "VALID", in_hw=phw, advance=False)
            self.pool2 = b.apool(1, 2, "VALID", in_hw=phw, advance=False)
            plain = dict(use_batch_norm=False, activation=None, bias=None, advance=False)
            self.conv1 = b.conv(filters // 2, 1, 1, "SAME", cin=pc, in_hw=self.pool1.out_hw, **plain)
            self.conv2 = b.conv(filters - filters // 2, 1, 1, "SAME", cin=pc, in_hw=self.pool1.out_hw, **plain)
            self.bn = b.batch_norm(activation=None, c=filters)
            assert self.pool1.out_hw == ip[1], (self.pool1.out_hw, ip)
        elif pc != filters:
            self.mode = "project"
            self.relu = Relu(b.cfg)
            self.conv = b.conv(filters, 1, 1, "SAME", cin=pc, in_hw=phw, use_batch_norm=True, activation=None, bias=None,
                               advance=False)
        self.out: Shape = (filters, ip[1])

    def forward(self, p):
        if self.mode == "same":
            return p
        p = self.relu(p)
        if self.mode == "project":
            return self.conv(p)
        p1 = self.conv1(self.pool1(p))
        shifted = torch.nn.functional.pad(p, (0, 1, 0, 1))[:, :, 1:, 1:]
        if p.is_contiguous(memory_format=torch.channels_last):
            shifted = shifted.contiguous(memory_format=torch.channels_last)
        p2 = self.conv2(self.pool2(shifted))
        return self.bn(ops.concat_channels([p1, p2]))


class Cell(nn.Module):
    def __init__(self, b: ConvNetBuilder, ip: Shape, p: Optional[Shape], filters: int, reduction: bool):
        super().__init__()
        self.cfg = b.cfg
        self.reduction = reduction
        b.push_scope("cell")
        self.adjust = Adjust(b, p, ip, filters)
        self.relu = Relu(b.cfg)
        self.base = b.conv(filters, 1, 1, "SAME", cin=ip[0], in_hw=ip[1], use_batch_norm=True, activation=None, bias=None,
                           advance=False)
        h: Shape = (filters, ip[1])
        pp = self.adjust.out
        sep = lambda inp, k, s=1: SepConv(b, inp, filters, k, s)
        pool = lambda kind, inp, s: (b.apool if kind == "avg" else b.mpool)(3, s, "SAME", in_hw=inp[1], advance=False)
        if not reduction:
            self.x1a, self.x1b = sep(h, 5), sep(pp, 3)
            self.x2a, self.x2b = sep(pp, 5), sep(pp, 3)
            self.x3 = pool("avg", h, 1)
            self.x4a, self.x4b = pool("avg", pp, 1), pool("avg", pp, 1)
            self.x5 = sep(h, 3)
            self.out: Shape = (6 * filters, ip[1])
        else:
            self.x1a, self.x1b = sep(h, 5, 2), sep(pp, 7, 2)
            self.x2a, self.x2b = pool("max", h, 2), sep(pp, 7, 2)
            self.x3a, self.x3b = pool("avg", h, 2), sep(pp, 5, 2)
            self.x4 = pool("avg", self.x1a.out, 1)
            self.x5a, self.x5b = sep(self.x1a.out, 3), pool("max", h, 2)
            self.out = (4 * filters, self.x1a.out[1])
        b.pop_scope()

    def forward(self, ip, p):
        add = lambda a, c: ops.add(a, c, self.cfg)
        p = self.adjust(ip if p is None else p)
        h = self.base(self.relu(ip))
        if not self.reduction:
            x1 = add(self.x1a(h), self.x1b(p))
            x2 = add(self.x2a(p), self.x2b(p))
            x3 = add(self.x3(h), p)
            x4 = add(self.x4a(p), self.x4b(p))
            x5 = add(self.x5(h), h)
            return ops.concat_channels([p, x1, x2, x3, x4, x5])
        x1 = add(self.x1a(h), self.x1b(p))
        x2 = add(self.x2a(h), self.x2b(p))
        x3 = add(self.x3a(h), self.x3b(p))
        x4 = add(x2, self.x4(x1))
        x5 = add(self.x5a(x1), self.x5b(h))
        return ops.concat_channels([x2, x3, x4, x5])


class NASNet(Model):
    default_learning_rate = 0.04

    def __init__(self, cfg: BuildConfig, name: str, image_size: int, n_cells: int, penultimate: int, stem_filters: int,
                 batch: int, keep_prob: float, skip_reduction: bool):
        cfg.use_batch_norm = True
        cfg.bn_decay, cfg.bn_eps, cfg.bn_scale = 0.9997, 1e-3, True
        super().__init__(cfg)
        self.name, self.image_size, self.default_batch_size = name, image_size, batch
        filters = penultimate // 24
        b = ConvNetBuilder(cfg, 3, (image_size, image_size))
        self.stem = b.conv(stem_filters, 3, 2, "VALID", use_batch_norm=True, activation=None, bias=None)
        cells = []
        ip: Shape = (stem_filters, b.hw)
        p: Optional[Shape] = None

        def push(f, reduction, keep_p=False):
            nonlocal ip, p
            cell = Cell(b, ip, p, f, reduction)
            cell.keep_p = keep_p      # the next cell's p stays this cell's p (the large model skips the reduction input)
            cells.append(cell)
            p, ip = (p if keep_p else ip), cell.out

        push(filters // 4, True)
        push(filters // 2, True)
        for stage, f in enumerate((filters, 2 * filters, 4 * filters)):
            if stage:
                push(f, True, keep_p=skip_reduction)
            for _ in range(n_cells):
                push(f, False)
        self.cells = nn.ModuleList(cells)
        self.relu = Relu(cfg)
        b.c, b.hw = ip
        self.gap = b.spatial_mean()
        self.drop = b.dropout(keep_prob)
        self.head = b.affine(cfg.num_classes, activation=None)

    def features(self, x):
        ip, p = self.stem(x), None
        for cell in self.cells:
            ip, p = cell(ip, p), (p if cell.keep_p else ip)
        return self.drop(self.gap(self.relu(ip)))


def nasnet_mobile(cfg):
    return NASNet(cfg, "nasnet", 224, 4, 1056, 32, 32, 0.5, skip_reduction=False)


def nasnet_large(cfg):
    return NASNet(cfg, "nasnetlarge", 331, 6, 4032, 96, 16, 0.5, skip_reduction=True)
