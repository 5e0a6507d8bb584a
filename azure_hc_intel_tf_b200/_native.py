"""ctypes access to the two in-tree native libraries (see build.py).

``host()`` returns libhvd_core.so (always available: it is built on first use with g++).
``cuda()`` returns libb200cnn.so, the sm_100a kernel library. On a machine with a GPU a missing
or unloadable kernel library is an error, never a silent fallback.
"""
from __future__ import annotations

import ctypes as C
import threading

from . import build as _build

_lock = threading.Lock()
_host = None
_cuda = None


class TensorDesc(C.Structure):
    _fields_ = [("numel", C.c_int64), ("elem_size", C.c_int32), ("dtype_id", C.c_int32),
                ("flags", C.c_int32), ("reserved", C.c_int32)]


class Segment(C.Structure):
    _fields_ = [("tensor", C.c_int32), ("bucket", C.c_int32), ("tensor_offset", C.c_int64),
                ("bucket_offset", C.c_int64), ("numel", C.c_int64)]


class Bucket(C.Structure):
    _fields_ = [("numel", C.c_int64), ("elem_size", C.c_int32), ("dtype_id", C.c_int32),
                ("first_segment", C.c_int32), ("num_segments", C.c_int32)]


def _declare_host(lib):
    vp, i32, i64, dbl, cp = C.c_void_p, C.c_int32, C.c_int64, C.c_double, C.c_char_p
    sig = {
        "hvd_plan_create": (vp, [C.POINTER(TensorDesc), i32, i64, i32]),
        "hvd_plan_create_tail": (vp, [C.POINTER(TensorDesc), i32, i64, i32, i64]),
        "hvd_plan_destroy": (None, [vp]),
        "hvd_plan_num_buckets": (i32, [vp]),
        "hvd_plan_num_segments": (i32, [vp]),
        "hvd_plan_get_bucket": (i32, [vp, i32, C.POINTER(Bucket)]),
        "hvd_plan_get_segment": (i32, [vp, i32, C.POINTER(Segment)]),
        "hvd_plan_hash": (C.c_uint64, [vp]),
        "hvd_tracker_create": (vp, [vp]),
        "hvd_tracker_destroy": (None, [vp]),
        "hvd_tracker_begin_step": (None, [vp]),
        "hvd_tracker_tensor_ready": (i32, [vp, i32, C.POINTER(i32), i32]),
        "hvd_tracker_pending": (i32, [vp]),
        "hvd_tracker_bucket_done": (None, [vp, i32]),
        "hvd_tracker_oldest_inflight_seconds": (dbl, [vp]),
        "hvd_timeline_open": (vp, [cp, i32]),
        "hvd_timeline_event": (None, [vp, cp, cp, i32, dbl, dbl]),
        "hvd_timeline_mark": (None, [vp, cp, i32, dbl]),
        "hvd_timeline_now_us": (dbl, [vp]),
        "hvd_timeline_close": (None, [vp]),
        "hvd_stall_start": (vp, [vp, dbl, dbl, dbl]),
        "hvd_stall_warnings": (i32, [vp]),
        "hvd_stall_stop": (None, [vp]),
        "hvd_launch_arith": (i32, [i32, i32, i32, i32, i32, C.POINTER(i32)]),
        "hvd_core_version": (cp, []),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return lib


def host():
    global _host
    with _lock:
        if _host is None:
            path = _build.build_host()
            _host = _declare_host(C.CDLL(str(path)))
    return _host


def cuda_available() -> bool:
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def cuda(build_if_missing: bool = True):
    """Load the sm_100a kernel library. Raises if it cannot be built or loaded."""
    global _cuda
    with _lock:
        if _cuda is None:
            path = _build.CUDA_LIB
            if build_if_missing:
                path = _build.build_cuda()
            elif not path.exists():
                raise RuntimeError(f"{path} is missing: run `python -m azure_hc_intel_tf_b200.build`")
            lib = C.CDLL(str(path), mode=C.RTLD_GLOBAL)
            from . import _cuda_sigs
            _cuda_sigs.declare(lib)
            _cuda = lib
    return _cuda


def cuda_loaded() -> bool:
    return _cuda is not None
