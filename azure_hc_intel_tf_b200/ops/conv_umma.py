"""Python face of the tcgen05/TMEM/TMA implicit-GEMM convolution kernels (csrc/conv). Filled in as the
kernels land; ``supports_*`` gate which shapes are routed here (ops/conv_impl.py)."""
from __future__ import annotations


def supports_fprop(x, w_lp, m) -> bool:
    return False


def supports_dgrad(dy, w_lp, m, x_shape) -> bool:
    return False


def supports_wgrad(x, dy, m) -> bool:
    return False


def supports_gemm(m_, n_, k_) -> bool:
    return False
