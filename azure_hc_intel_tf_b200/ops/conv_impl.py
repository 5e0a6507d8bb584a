"""Convolution and GEMM back ends for the b200 ops layer.

``B200_CONV`` selects the implementation:

* ``umma``  — the hand-written tcgen05/TMEM/TMA implicit-GEMM kernels (csrc/conv); shapes the kernels do
  not cover raise;
* ``auto`` (default) — ``umma`` where the kernels cover the shape, otherwise the library fallback below;
* ``cudnn`` — library fallback everywhere (cuDNN through ``aten::convolution``, cuBLAS through ``mm``).

``path_counts`` records how many calls took which path so a run can report what actually executed.

Counterpart in the reference stack: MKL-DNN convolution forward / backward-data / backward-weights and
inner-product primitives (K1–K3, K8 in SURVEY.md §2.6).
"""
from __future__ import annotations

import os
from collections import Counter

import torch

path_counts = Counter()
_BF16 = torch.bfloat16


def mode() -> str:
    return os.environ.get("B200_CONV", "auto")


def _umma():
    from . import conv_umma
    return conv_umma


def _sym(m):
    (pt, pb), (pl, pr) = m.pad
    return pt == pb and pl == pr


def _groups(m) -> int:
    return getattr(m, "groups", 1)


def _padded(x, m):
    (pt, pb), (pl, pr) = m.pad
    return torch.nn.functional.pad(x, (pl, pr, pt, pb))


# ------------------------------------------------------------------------------------ conv

def fprop(x, w_lp, m, want_stats: bool = False):
    """x: bf16 NHWC-storage [N,C,H,W]; w_lp: bf16 [K,C,R,S] channels_last (KRSC storage). → y bf16.
    With ``want_stats`` returns ``(y, partials)`` where partials is [rows, 2, K] fp32 (per-CTA Σy, Σy² per
    channel, produced by the conv epilogue) or None when the path taken cannot produce them."""
    md = mode()
    if _groups(m) > 1:
        if md != "cudnn" and _umma().supports_depthwise(x.shape, m):
            path_counts["fprop_depthwise"] += 1
            y = _umma().dw_fprop(x, w_lp, m)
        else:
            if md == "umma":
                raise RuntimeError(f"B200_CONV=umma but the depthwise kernels do not cover {m.extra_repr()}")
            path_counts["fprop_lib"] += 1
            y = _fprop_lib(x, w_lp, m)
        return (y, None) if want_stats else y
    if md != "cudnn":
        u = _umma()
        if u.supports_fprop(x, w_lp, m):
            path_counts["fprop_umma"] += 1
            if want_stats:
                return u.fprop(x, w_lp, m, want_stats=True)     # (y, [rows,2,K] partial sums)
            return u.fprop(x, w_lp, m)
        if md == "umma":
            raise RuntimeError(f"B200_CONV=umma but fprop does not cover {m.extra_repr()}")
    path_counts["fprop_lib"] += 1
    y = _fprop_lib(x, w_lp, m)
    return (y, None) if want_stats else y


def _fprop_lib(x, w_lp, m):
    if _sym(m):
        y = torch.ops.aten.convolution(x, w_lp, None, m.stride, (m.pad[0][0], m.pad[1][0]), (1, 1), False, (0, 0), _groups(m))
    else:
        y = torch.ops.aten.convolution(_padded(x, m), w_lp, None, m.stride, (0, 0), (1, 1), False, (0, 0), _groups(m))
    return y if y.is_contiguous(memory_format=torch.channels_last) else y.contiguous(memory_format=torch.channels_last)


def dgrad(dy, w_lp, m, x_shape):
    md = mode()
    if _groups(m) > 1 and md != "cudnn" and _umma().supports_depthwise(x_shape, m):
        path_counts["dgrad_depthwise"] += 1
        return _umma().dw_dgrad(dy, w_lp, m, x_shape)
    if md != "cudnn" and _groups(m) == 1:
        u = _umma()
        if u.supports_dgrad(dy, w_lp, m, x_shape):
            path_counts["dgrad_umma"] += 1
            return u.dgrad(dy, w_lp, m, x_shape)
        if md == "umma":
            raise RuntimeError(f"B200_CONV=umma but dgrad does not cover {m.extra_repr()}")
    path_counts["dgrad_lib"] += 1
    n, c, h, w = x_shape
    (pt, pb), (pl, pr) = m.pad
    if _sym(m):
        xs, pad = (n, c, h, w), (pt, pl)
    else:
        xs, pad = (n, c, h + pt + pb, w + pl + pr), (0, 0)
    fake_x = dy.new_empty(xs).contiguous(memory_format=torch.channels_last)   # only its shape is used
    dx, _, _ = torch.ops.aten.convolution_backward(dy, fake_x, w_lp, None, m.stride, pad, (1, 1), False, (0, 0), _groups(m),
                                                   [True, False, False])
    if not _sym(m):
        dx = dx[:, :, pt:pt + h, pl:pl + w]
    return dx.contiguous(memory_format=torch.channels_last)


def wgrad(x, dy, m, out=None):
    """Returns fp32 [K,C,R,S] (channels_last); when ``out`` (fp32 view, same layout) is given it is filled."""
    md = mode()
    if _groups(m) > 1 and md != "cudnn" and _umma().supports_depthwise(x.shape, m):
        path_counts["wgrad_depthwise"] += 1
        return _umma().dw_wgrad(x, dy, m, out)
    if md != "cudnn" and _groups(m) == 1:
        u = _umma()
        if u.supports_wgrad(x, dy, m):
            path_counts["wgrad_umma"] += 1
            return u.wgrad(x, dy, m, out)
        if md == "umma":
            raise RuntimeError(f"B200_CONV=umma but wgrad does not cover {m.extra_repr()}")
    path_counts["wgrad_lib"] += 1
    k, c = m.cout, m.cin // _groups(m)
    r, s = m.k
    xx, pad = (x, (m.pad[0][0], m.pad[1][0])) if _sym(m) else (_padded(x, m), (0, 0))
    fake_w = x.new_empty((k, c, r, s)).contiguous(memory_format=torch.channels_last)
    _, dw, _ = torch.ops.aten.convolution_backward(dy, xx, fake_w, None, m.stride, pad, (1, 1), False, (0, 0), _groups(m),
                                                   [False, True, False])
    if out is not None:
        out.copy_(dw)
        return out
    return dw.float()


# ------------------------------------------------------------------------------------ first layer (Cin = 3)

def first_layer(x, m) -> bool:
    """True when the small-Cin path (explicit patch matrix + tensor-core GEMM) handles this layer."""
    return mode() != "cudnn" and m.cin < 8 and not x.requires_grad and _umma().supports_first_layer(x, m)


def first_fprop(x, w_lp, m, want_stats):
    path_counts["fprop_umma_first"] += 1
    return _umma().first_layer_fprop(x, w_lp, m, want_stats)


def first_wgrad(patches, dy, m, out=None):
    path_counts["wgrad_umma_first"] += 1
    return _umma().first_layer_wgrad(patches, dy, m, out)


# ------------------------------------------------------------------------------------ dense GEMMs

def gemm_nt(a, b, out_fp32: bool = False):
    """a[M,K] · b[N,K]ᵀ → [M,N] bf16 (or fp32)."""
    md = mode()
    if md != "cudnn":
        u = _umma()
        if u.supports_gemm(a.shape[0], b.shape[0], a.shape[1]):
            path_counts["gemm_umma"] += 1
            return u.gemm_nt(a, b, out_fp32=out_fp32)
    path_counts["gemm_lib"] += 1
    r = torch.mm(a, b.t())
    return r.float() if out_fp32 else r


def gemm_nn(a, b, k=None):
    """a[M,K(ld)] · b[K,N] → bf16 [M,N]. ``k``: number of valid columns of a when it is zero-padded."""
    md = mode()
    kk = b.shape[0]
    if md != "cudnn" and b.shape[1] % 64 == 0 and a.shape[1] % 8 == 0:
        u = _umma()
        if u.supports_gemm(a.shape[0], b.shape[1], 64):
            path_counts["gemm_umma"] += 1
            return u.gemm_nn(a, b, k=kk)
    path_counts["gemm_lib"] += 1
    return torch.mm(a[:, :kk], b)


def gemm_tn_f32(a, b, out=None, out_features=None):
    """a[K,M(ld)]ᵀ · b[K,N] → fp32 [M,N] (weight gradient of a dense layer); a may be column-padded."""
    md = mode()
    of = out_features if out_features is not None else a.shape[1]
    if md != "cudnn" and b.shape[1] % 64 == 0 and a.shape[1] % 8 == 0:
        u = _umma()
        if u.supports_gemm(of, b.shape[1], 64):
            path_counts["gemm_umma"] += 1
            o = out if (out is not None and out.is_contiguous()) else None
            r = u.dense_wgrad(b, a, of, out=o)
            if out is not None and o is None:
                out.copy_(r)
                return out
            return r
    path_counts["gemm_lib"] += 1
    a = a[:, :of]
    r = torch.mm(a.t(), b)
    if out is not None:
        out.copy_(r)
        return out
    return r.float()
