"""Synthetic ImageNet input (K11 in SURVEY.md §2.6).

tf_cnn_benchmarks falls back to synthetic data when ``--data_dir`` is absent: images drawn once from a
truncated normal (mean 127, stddev 60), labels uniform, the same batch reused every step. The reference
points ``--data_dir`` at a TFRecord subset (benchmark-scripts/run-tf-sing-ucx-openmpi.sh:19,80); that
dataset does not exist on this box, so only the synthetic source is provided.

Two sources:

* :class:`DeviceSynthetic` — the batch lives on the device (tf_cnn_benchmarks' synthetic mode).
* :class:`HostSynthetic` — the batch is uint8 NHWC in pinned host memory (what a JPEG decoder hands over),
  copied host→device every step on a copy stream, double-buffered, then cast/normalised on the device.
  This is the source the end-to-end number is measured with.

Both yield ``(images, labels)`` with images as a logical-NCHW ``channels_last`` tensor in the compute dtype,
normalised to roughly [-1, 1] (``x/127.5 - 1``, tf_cnn_benchmarks' ``normalized_image``).
"""
from __future__ import annotations

import torch


def _truncated_normal_u8(shape, gen, mean=127.0, std=60.0):
    x = torch.empty(shape, dtype=torch.float32)
    torch.nn.init.trunc_normal_(x, mean=mean, std=std, a=mean - 2 * std, b=mean + 2 * std, generator=gen)
    return x.round_().clamp_(0, 255).to(torch.uint8)


class DeviceSynthetic:
    def __init__(self, batch: int, image_size: int, num_classes: int, device, dtype, seed: int = 0,
                 own_kernels: bool = True):
        gen = torch.Generator().manual_seed(seed)
        u8 = _truncated_normal_u8((batch, image_size, image_size, 3), gen)
        img = (u8.float() / 127.5 - 1.0).to(dtype)
        self.images = img.to(device).permute(0, 3, 1, 2)   # NHWC storage, NCHW logical = channels_last
        self.labels = torch.randint(0, num_classes - 1, (batch,), generator=gen).to(device)
        self.h2d_bytes = 0

    def next(self):
        return self.images, self.labels


class HostSynthetic:
    """Pinned-host uint8 source with a double-buffered async H2D copy and an on-device normalise/cast."""

    def __init__(self, batch: int, image_size: int, num_classes: int, device, dtype, seed: int = 0, nbuf: int = 2,
                 own_kernels: bool = True):
        self.own_kernels = own_kernels
        gen = torch.Generator().manual_seed(seed)
        self.device, self.dtype = torch.device(device), dtype
        cuda = self.device.type == "cuda"
        self.host_images = _truncated_normal_u8((batch, image_size, image_size, 3), gen)
        self.host_labels = torch.randint(0, num_classes - 1, (batch,), generator=gen, dtype=torch.int64)
        if cuda:
            self.host_images = self.host_images.pin_memory()
            self.host_labels = self.host_labels.pin_memory()
        self.h2d_bytes = self.host_images.numel() + self.host_labels.numel() * 8
        self.nbuf = nbuf
        self.stream = torch.cuda.Stream() if cuda else None
        self.dev_u8 = [torch.empty_like(self.host_images, device=self.device) for _ in range(nbuf)]
        self.dev_lab = [torch.empty_like(self.host_labels, device=self.device) for _ in range(nbuf)]
        self.ready = [None] * nbuf
        self.consumed = [None] * nbuf
        self.i = 0
        self._issue(0)

    def _issue(self, slot):
        if self.stream is None:
            self.dev_u8[slot].copy_(self.host_images)
            self.dev_lab[slot].copy_(self.host_labels)
            return
        if self.consumed[slot] is not None:
            self.stream.wait_event(self.consumed[slot])
        with torch.cuda.stream(self.stream):
            self.dev_u8[slot].copy_(self.host_images, non_blocking=True)
            self.dev_lab[slot].copy_(self.host_labels, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.stream)
        self.ready[slot] = ev

    def next(self):
        slot = self.i % self.nbuf
        nxt = (self.i + 1) % self.nbuf
        self.i += 1
        if self.stream is not None:
            torch.cuda.current_stream().wait_event(self.ready[slot])
        u8, lab = self.dev_u8[slot], self.dev_lab[slot]
        img = normalize_u8(u8, self.dtype, self.own_kernels)
        labels = lab.clone()
        if self.stream is not None:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream())
            self.consumed[slot] = ev
        self._issue(nxt)     # prefetch the next step's batch while this step computes
        return img, labels


def normalize_u8(u8_nhwc: torch.Tensor, dtype, own_kernels: bool = True) -> torch.Tensor:
    """uint8 NHWC → normalised ``dtype``, returned as a logical-NCHW channels_last tensor."""
    if u8_nhwc.is_cuda and own_kernels and dtype == torch.bfloat16:
        from . import b200
        return b200.normalize_u8(u8_nhwc)
    return (u8_nhwc.to(torch.float32) / 127.5 - 1.0).to(dtype).permute(0, 3, 1, 2)
