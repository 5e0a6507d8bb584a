"""Python face of the tcgen05/TMEM/TMA implicit-GEMM convolution kernels (csrc/conv/conv_gemm.cu,
conv_wgrad.cu; C ABI in include/b200conv.h). ``supports_*`` ask the library which geometries the kernels
cover; ops/conv_impl.py routes everything else to the library fallback.

Reference role: MKL-DNN convolution primitives under --mkl=TRUE (run-tf-sing-ucx-openmpi.sh:76), K1–K3."""
from __future__ import annotations

import ctypes as C
import os

import torch

from .._cuda_sigs import ConvGeom
from . import b200
from .b200 import _check, _ptr, _stream, empty_nhwc, lib

FPROP, DGRAD, WGRAD = 0, 1, 2
_BF16 = torch.bfloat16
_disabled = set(filter(None, os.environ.get("B200_CONV_DISABLE", "").split(",")))   # e.g. "wgrad,dgrad"


def geom(x_shape, m) -> ConvGeom:
    n, c, h, w = x_shape
    p, q = m.out_hw
    (pt, pb), (pl, pr) = m.pad
    return ConvGeom(n, h, w, c, m.cout, m.k[0], m.k[1], p, q, m.stride[0], m.stride[1], pt, pl, pb, pr)


def _ok(g, op, name) -> bool:
    if name in _disabled:
        return False
    fn = getattr(lib(), "b200_conv_supported", None)
    return bool(fn and fn(C.byref(g), op))


def supports_fprop(x, w_lp, m) -> bool:
    return _ok(geom(x.shape, m), FPROP, "fprop")


def supports_dgrad(dy, w_lp, m, x_shape) -> bool:
    return _ok(geom(x_shape, m), DGRAD, "dgrad")


def supports_wgrad(x, dy, m) -> bool:
    return _ok(geom(x.shape, m), WGRAD, "wgrad")


def supports_gemm(m_, n_, k_) -> bool:
    return "gemm" not in _disabled and k_ % 8 == 0


def fprop(x, w_lp, m, want_stats: bool = False):
    g = geom(x.shape, m)
    y = empty_nhwc(g.n, g.k, g.p, g.q, x.device)
    stats = None
    if want_stats:
        rows = lib().b200_conv_stats_rows(C.byref(g))
        stats = torch.empty((rows, 2, g.k), dtype=torch.float32, device=x.device)
    _check(lib().b200_conv_fprop(_ptr(x), _ptr(w_lp), _ptr(y), C.byref(g), _ptr(stats), _stream()), "conv_fprop")
    if want_stats:   # the leading rows are kernel scratch; results are one row per CTA after them
        return y, stats[lib().b200_conv_stats_first_result_row(C.byref(g)):]
    return y


def stats_reduce(stats):
    rows, _, k = stats.shape
    sums = torch.empty((2, k), dtype=torch.float32, device=stats.device)
    _check(lib().b200_conv_stats_reduce(_ptr(stats), rows, k, _ptr(sums), _stream()), "conv_stats_reduce")
    return sums


_ws = {}


def dgrad(dy, w_lp, m, x_shape):
    g = geom(x_shape, m)
    dx = empty_nhwc(g.n, g.c, g.h, g.w, dy.device)
    if lib().b200_conv_supported(C.byref(g), DGRAD) == 2:
        # general strided filter: one stride-1 launch per stride phase (csrc/conv/conv_gemm.cu b200_conv_dgrad_strided)
        ws = _workspace(lib().b200_conv_dgrad_strided_workspace_bytes(C.byref(g)), dy.device)
        _check(lib().b200_conv_dgrad_strided(_ptr(dy), _ptr(w_lp), _ptr(dx), C.byref(g), _ptr(ws), _stream()),
               "conv_dgrad_strided")
        b200.count_launch(2 * m.stride[0] * m.stride[1] - 1)
        return dx
    _check(lib().b200_conv_dgrad(_ptr(dy), _ptr(w_lp), _ptr(dx), C.byref(g), _stream()), "conv_dgrad")
    return dx


def _workspace(need, device):
    """Split-K partial-sum workspace, one per device and stream (kernels using it are ordered on that stream)."""
    if need <= 0:
        return None
    key = (device.index, torch.cuda.current_stream().cuda_stream)
    ws = _ws.get(key)
    if ws is None or ws.numel() < need:
        ws = torch.empty(max(need, 100 << 20), dtype=torch.uint8, device=device)
        _ws[key] = ws
    return ws


def wgrad(x, dy, m, out=None):
    g = geom(x.shape, m)
    if out is None:
        out = torch.empty((g.k, g.r, g.s, g.c), dtype=torch.float32, device=x.device).permute(0, 3, 1, 2)
    ws = _workspace(lib().b200_conv_wgrad_workspace_bytes(C.byref(g)), x.device)
    _check(lib().b200_conv_wgrad(_ptr(x), _ptr(dy), _ptr(out), C.byref(g), _ptr(ws), _stream()), "conv_wgrad")
    b200.count_launch(1 if ws is not None else 0)
    return out


# ------------------------------------------------------------------------------------ depthwise (csrc/kernels/depthwise.cu)

def supports_depthwise(x_shape, m) -> bool:
    if "depthwise" in _disabled or m.groups != m.cin or m.cin != m.cout:
        return False
    fn = getattr(lib(), "b200_dwconv_supported", None)
    return bool(fn and fn(C.byref(geom(x_shape, m))))


def dw_fprop(x, w_lp, m):
    g = geom(x.shape, m)
    y = empty_nhwc(g.n, g.k, g.p, g.q, x.device)
    _check(lib().b200_dwconv_fprop(_ptr(x), _ptr(w_lp), _ptr(y), C.byref(g), _stream()), "dwconv_fprop")
    return y


def dw_dgrad(dy, w_lp, m, x_shape):
    g = geom(x_shape, m)
    dx = empty_nhwc(g.n, g.c, g.h, g.w, dy.device)
    _check(lib().b200_dwconv_dgrad(_ptr(dy), _ptr(w_lp), _ptr(dx), C.byref(g), _stream()), "dwconv_dgrad")
    return dx


def dw_wgrad(x, dy, m, out=None):
    g = geom(x.shape, m)
    if out is None:
        out = torch.empty((g.k, g.r, g.s, 1), dtype=torch.float32, device=x.device).permute(0, 3, 1, 2)
    ws = _workspace(lib().b200_dwconv_wgrad_workspace_bytes(C.byref(g)), x.device)
    _check(lib().b200_dwconv_wgrad(_ptr(x), _ptr(dy), _ptr(out), C.byref(g), _ptr(ws), _stream()), "dwconv_wgrad")
    b200.count_launch(1)
    return out


def gemm_nt(a, b, out_fp32: bool = False):
    mm, kk = a.shape
    nn = b.shape[0]
    out = torch.empty((mm, nn), dtype=torch.float32 if out_fp32 else _BF16, device=a.device)
    _check(lib().b200_gemm_nt(_ptr(a), _ptr(b), _ptr(out), mm, nn, kk, nn, int(out_fp32), _stream()), "gemm_nt")
    return out


def gemm_nn(a, b, k=None):
    """a[M, lda] (only the first k columns are used; the rest must be zero) · b[k, N] → bf16 [M, N]."""
    mm, lda = a.shape
    kk, nn = b.shape
    out = torch.empty((mm, nn), dtype=_BF16, device=a.device)
    _check(lib().b200_gemm_nn(_ptr(a), lda, _ptr(b), _ptr(out), mm, nn, kk if k is None else k, _stream()), "gemm_nn")
    return out


def dense_wgrad(x, dy, out_features, out=None):
    """dw[out, in] fp32 = dy[B, ld]ᵀ · x[B, in]; dy may be padded to ld >= out_features columns."""
    bsz, infe = x.shape
    ld = dy.shape[1]
    if out is None:
        out = torch.empty((out_features, infe), dtype=torch.float32, device=x.device)
    ws = _workspace(lib().b200_dense_wgrad_workspace_bytes(bsz, infe, out_features), x.device)
    _check(lib().b200_dense_wgrad(_ptr(x), _ptr(dy), _ptr(out), bsz, infe, out_features, ld, _ptr(ws), _stream()),
           "dense_wgrad")
    b200.count_launch(1 if ws is not None else 0)
    return out


# ------------------------------------------------------------------------------------ first layer (Cin = 3)

def supports_first_layer(x, m) -> bool:
    return "first" not in _disabled and m.cin < 8 and m.cout % 8 == 0


def first_layer_kpad(m) -> int:
    """Row length of the patch matrix: R*S*Cin rounded up to the 8-element (16-byte) rows TMA needs - 152 for the 7x7x3
    ImageNet stem; the GEMM's last 64-wide k-block is zero-filled by TMA past it. B200_FIRST_KPAD=64 restores rounding
    to whole k-blocks (192: 26 % more patch-matrix bytes written once and read twice per step)."""
    unit = 64 if os.environ.get("B200_FIRST_KPAD") == "64" else 8
    return (m.k[0] * m.k[1] * m.cin + unit - 1) // unit * unit


def first_layer_fprop(x, w_lp, m, want_stats: bool):
    """Explicit im2col (patch matrix A) + tcgen05 GEMM. Returns (y, partial stats or None, A); A is kept for wgrad."""
    g = geom(x.shape, m)
    kdim, kpad = m.k[0] * m.k[1] * m.cin, first_layer_kpad(m)
    rows = g.n * g.p * g.q
    a = torch.empty((rows, kpad), dtype=_BF16, device=x.device)
    _check(lib().b200_im2col_small_c(_ptr(x), _ptr(a), C.byref(g), kpad, _stream()), "im2col_small_c")
    # filter [K, C, R, S] channels_last = KRSC storage = [K, R*S*C] row-major; pad the rows to kpad
    w2 = torch.zeros((m.cout, kpad), dtype=_BF16, device=x.device)
    w2[:, :kdim] = w_lp.permute(0, 2, 3, 1).reshape(m.cout, kdim)
    y = empty_nhwc(g.n, g.k, g.p, g.q, x.device)
    stats = None
    if want_stats:
        srows = lib().b200_gemm_stats_rows(rows, g.k)
        stats = torch.empty((srows, 2, g.k), dtype=torch.float32, device=x.device)
    _check(lib().b200_gemm_nt_stats(_ptr(a), _ptr(w2), _ptr(y), rows, g.k, kpad, g.k, 0, _ptr(stats), _stream()),
           "gemm_nt_stats")
    if stats is not None:
        stats = stats[lib().b200_gemm_stats_first_result_row(rows, g.k):]
    return y, stats, a


def first_layer_wgrad(a, dy, m, out=None):
    """dw[K, C, R, S] fp32 from the saved patch matrix: dense_wgrad(A, dY) then drop the K padding."""
    kdim = m.k[0] * m.k[1] * m.cin
    n, k, p, q = dy.shape
    dy2 = dy.permute(0, 2, 3, 1).reshape(n * p * q, k)
    full = dense_wgrad(a, dy2, k)                                  # [K, kpad] fp32
    dw = full[:, :kdim].reshape(k, m.k[0], m.k[1], m.cin).permute(0, 3, 1, 2)
    if out is not None:
        out.copy_(dw)
        return out
    return dw
