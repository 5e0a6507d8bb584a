"""Gradient compression, Horovod-style (``hvd.Compression.none`` / ``.fp16``). The reference does not
enable compression (no flag at run-tf-sing-ucx-openmpi.sh:62-81); on B200 the 16-bit wire format is bf16."""
from __future__ import annotations

import torch


class _NoneCompressor:
    @staticmethod
    def compress(tensor):
        return tensor, None

    @staticmethod
    def decompress(tensor, ctx):
        return tensor


class _BF16Compressor:
    @staticmethod
    def compress(tensor):
        if tensor.dtype.is_floating_point and tensor.dtype != torch.bfloat16:
            return tensor.to(torch.bfloat16), tensor.dtype
        return tensor, None

    @staticmethod
    def decompress(tensor, ctx):
        return tensor.to(ctx) if ctx is not None else tensor


class Compression:
    none = _NoneCompressor
    fp16 = _BF16Compressor   # Horovod's name; bf16 on the wire here
    bf16 = _BF16Compressor
