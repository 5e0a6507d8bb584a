"""ctypes signatures of libb200cnn.so (mirrors include/b200cnn.h, include/b200comm.h, include/b200conv.h)."""
from __future__ import annotations

import ctypes as C

vp, i32, i64, f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_float


class PoolGeom(C.Structure):
    _fields_ = [(n, i32) for n in ("n", "h", "w", "c", "p", "q", "kh", "kw", "sh", "sw", "pad_t", "pad_l")]


class ConvGeom(C.Structure):
    _fields_ = [(n, i32) for n in ("n", "h", "w", "c", "k", "r", "s", "p", "q", "stride_h", "stride_w",
                                   "pad_t", "pad_l", "pad_b", "pad_r")]


SIGS = {
    "b200_version": (C.c_char_p, []),
    "b200_last_error": (C.c_char_p, []),
    "b200_sgd_momentum_flat": (i32, [vp, vp, vp, vp, i64, vp, vp, vp, vp, f32, f32, f32, f32, vp, vp]),
    "b200_rmsprop_flat": (i32, [vp, vp, vp, vp, vp, i64, vp, vp, f32, f32, f32, f32, f32, f32, vp, vp]),
    "b200_adam_flat": (i32, [vp, vp, vp, vp, vp, i64, vp, vp, f32, f32, f32, f32, f32, f32, vp, vp]),
    "b200_lars_norms": (i32, [vp, vp, i64, vp, vp, f32, vp]),
    "b200_lars_trust_det": (i32, [vp, vp, i64, vp, vp, vp, vp, vp, vp, i32, f32, f32, f32, f32, vp]),
    "b200_lars_trust": (i32, [vp, vp, vp, vp, i32, f32, f32, f32, vp]),
    "b200_bn_stats_scratch_bytes": (i64, [i64, i32]),
    "b200_bn_stats": (i32, [vp, i64, i32, vp, vp, vp, vp, f32, f32, vp, vp, vp, vp, vp, vp]),
    "b200_bn_finalize": (i32, [vp, i64, i32, vp, vp, vp, vp, f32, f32, vp, vp, vp, vp, vp]),
    "b200_bn_finalize_partials": (i32, [vp, i64, i64, i32, vp, vp, vp, vp, f32, f32, vp, vp, vp, vp, vp]),
    "b200_bn_apply": (i32, [vp, vp, vp, vp, i64, i32, vp, vp, i32, vp]),
    "b200_bn_bwd_reduce": (i32, [vp, vp, vp, vp, i64, i32, vp, vp, vp, vp, i32, vp, vp, vp, vp]),
    "b200_bn_bwd_apply": (i32, [vp, vp, vp, vp, vp, vp, i64, i32, vp, vp, vp, vp, vp, vp, i32, vp]),
    "b200_bias_act": (i32, [vp, vp, vp, i64, i32, i32, vp]),
    "b200_bias_act_bwd": (i32, [vp, vp, vp, vp, i64, i32, i32, vp, vp]),
    "b200_maxpool_fwd": (i32, [vp, vp, vp, C.POINTER(PoolGeom), vp]),
    "b200_maxpool_bwd": (i32, [vp, vp, vp, C.POINTER(PoolGeom), vp]),
    "b200_avgpool_fwd": (i32, [vp, vp, C.POINTER(PoolGeom), vp]),
    "b200_avgpool_bwd": (i32, [vp, vp, C.POINTER(PoolGeom), vp]),
    "b200_spatial_mean_fwd": (i32, [vp, vp, i32, i32, i32, vp]),
    "b200_spatial_mean_bwd": (i32, [vp, vp, i32, i32, i32, vp]),
    "b200_softmax_xent": (i32, [vp, i32, vp, i32, i32, vp, vp, vp, i32, f32, vp]),
    "b200_normalize_u8": (i32, [vp, vp, i64, i32, i32, vp]),
    "b200_add": (i32, [vp, vp, vp, i64, vp]),
    "b200_add_masked": (i32, [vp, vp, vp, vp, i64, vp]),
    "b200_add_act": (i32, [vp, vp, vp, i64, i32, vp]),
    "b200_relu_bwd": (i32, [vp, vp, vp, i64, vp]),
    "b200_cast_f32_bf16": (i32, [vp, vp, i64, vp]),
}

# conv / GEMM era entry points (kept in a second table only for readability)
OPTIONAL = {
    "b200_bn_stats_noalloc": (i32, [vp, i64, i32, vp, vp, vp, vp, f32, f32, vp, vp, vp, vp, vp]),
    "b200_bn_stats_noalloc_max_bytes": (i64, []),
    "b200_conv_supported": (i32, [C.POINTER(ConvGeom), i32]),
    "b200_conv_min_channels": (i32, []),
    "b200_conv_fprop": (i32, [vp, vp, vp, C.POINTER(ConvGeom), vp, vp]),
    "b200_conv_stats_rows": (i64, [C.POINTER(ConvGeom)]),
    "b200_conv_stats_first_result_row": (i64, [C.POINTER(ConvGeom)]),
    "b200_gemm_stats_first_result_row": (i64, [i32, i32]),
    "b200_conv_dgrad": (i32, [vp, vp, vp, C.POINTER(ConvGeom), vp]),
    "b200_conv_dgrad_strided": (i32, [vp, vp, vp, C.POINTER(ConvGeom), vp, vp]),
    "b200_conv_dgrad_strided_workspace_bytes": (i64, [C.POINTER(ConvGeom)]),
    "b200_conv_wgrad": (i32, [vp, vp, vp, C.POINTER(ConvGeom), vp, vp]),
    "b200_conv_wgrad_workspace_bytes": (i64, [C.POINTER(ConvGeom)]),
    "b200_dwconv_supported": (i32, [C.POINTER(ConvGeom)]),
    "b200_dwconv_fprop": (i32, [vp, vp, vp, C.POINTER(ConvGeom), vp]),
    "b200_dwconv_dgrad": (i32, [vp, vp, vp, C.POINTER(ConvGeom), vp]),
    "b200_dwconv_wgrad": (i32, [vp, vp, vp, C.POINTER(ConvGeom), vp, vp]),
    "b200_dwconv_wgrad_workspace_bytes": (i64, [C.POINTER(ConvGeom)]),
    "b200_gemm_nt": (i32, [vp, vp, vp, i32, i32, i32, i32, i32, vp]),
    "b200_gemm_nt_stats": (i32, [vp, vp, vp, i32, i32, i32, i32, i32, vp, vp]),
    "b200_gemm_stats_rows": (i64, [i32, i32]),
    "b200_im2col_small_c": (i32, [vp, vp, C.POINTER(ConvGeom), i32, vp]),
    "b200_gemm_nn": (i32, [vp, i32, vp, vp, i32, i32, i32, vp]),
    "b200_dense_wgrad": (i32, [vp, vp, vp, i32, i32, i32, i32, vp, vp]),
    "b200_dense_wgrad_workspace_bytes": (i64, [i32, i32, i32]),
    "b200_conv_stats_reduce": (i32, [vp, i64, i32, vp, vp]),
}


def declare(lib):
    for name, (res, args) in SIGS.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    # every entry point is required: a library that lacks one is stale or mis-built, and silently routing its callers
    # to a fallback would hide that (tests/test_abi.py checks headers, signatures and exports against each other)
    for name, (res, args) in OPTIONAL.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return lib
