"""Python face of the C++ static fusion planner (csrc/core/hvd_core.cpp, include/hvd_core.h).

Horovod's coordinator negotiates a fused response list every cycle and caps each fusion buffer at
HOROVOD_FUSION_THRESHOLD (run-tf-sing-ucx-openmpi.sh:105). A CNN's gradient order is static, so the plan
is built once here; ranks compare ``plan.hash`` at the first step.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import List, Sequence

from .. import _native

FLAG_DECAY = 1
FLAG_LARS = 2
DEFAULT_THRESHOLD = 134217728
ALIGN_ELEMS = 1024   # every segment starts on a 4 KiB (fp32) boundary: vector loads, TMA and multimem safe


@dataclass
class Seg:
    tensor: int
    bucket: int
    tensor_offset: int
    bucket_offset: int
    numel: int


@dataclass
class Buck:
    numel: int
    elem_size: int
    dtype_id: int
    first_segment: int
    num_segments: int
    start: int = 0   # element offset of this bucket in the flat buffer of its dtype


class FusionPlan:
    def __init__(self, numels: Sequence[int], elem_sizes: Sequence[int], dtype_ids: Sequence[int],
                 flags: Sequence[int], threshold_bytes: int = DEFAULT_THRESHOLD, align_elems: int = ALIGN_ELEMS,
                 tail_bytes: int = 0):
        lib = _native.host()
        n = len(numels)
        arr = (_native.TensorDesc * max(n, 1))()
        for i in range(n):
            arr[i].numel, arr[i].elem_size, arr[i].dtype_id, arr[i].flags = \
                int(numels[i]), int(elem_sizes[i]), int(dtype_ids[i]), int(flags[i])
        self._lib = lib
        self._h = lib.hvd_plan_create_tail(arr, n, int(threshold_bytes), int(align_elems), int(tail_bytes))
        if not self._h:
            raise ValueError("fusion planner rejected its input")
        self.num_tensors = n
        self.threshold_bytes = threshold_bytes
        self.buckets: List[Buck] = []
        self.segments: List[Seg] = []
        b, s = _native.Bucket(), _native.Segment()
        for i in range(lib.hvd_plan_num_buckets(self._h)):
            lib.hvd_plan_get_bucket(self._h, i, C.byref(b))
            self.buckets.append(Buck(b.numel, b.elem_size, b.dtype_id, b.first_segment, b.num_segments))
        for i in range(lib.hvd_plan_num_segments(self._h)):
            lib.hvd_plan_get_segment(self._h, i, C.byref(s))
            self.segments.append(Seg(s.tensor, s.bucket, s.tensor_offset, s.bucket_offset, s.numel))
        self.hash = int(lib.hvd_plan_hash(self._h))
        # flat layout: buckets of one dtype are laid out back to back, so a tensor split over
        # consecutive buckets is still one contiguous range
        run = {}
        for bk in self.buckets:
            bk.start = run.get(bk.dtype_id, 0)
            run[bk.dtype_id] = bk.start + bk.numel
        self.total_elems = run

    def tensor_range(self, t: int):
        """(dtype_id, start, numel) of tensor t in the flat buffer."""
        segs = [s for s in self.segments if s.tensor == t]
        first = segs[0]
        bk = self.buckets[first.bucket]
        start = bk.start + first.bucket_offset
        pos = start
        for s in segs:
            assert self.buckets[s.bucket].start + s.bucket_offset == pos, "tensor is not contiguous in the flat layout"
            pos += s.numel
        return bk.dtype_id, start, pos - start

    def tracker(self):
        return Tracker(self)

    def close(self):
        if self._h:
            self._lib.hvd_plan_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Tracker:
    def __init__(self, plan: FusionPlan):
        self._lib = plan._lib
        self._plan = plan
        self._h = self._lib.hvd_tracker_create(plan._h)
        self._out = (C.c_int32 * max(len(plan.buckets), 1))()

    def begin_step(self):
        self._lib.hvd_tracker_begin_step(self._h)

    def ready(self, tensor: int) -> List[int]:
        n = self._lib.hvd_tracker_tensor_ready(self._h, tensor, self._out, len(self._out))
        if n < 0:
            raise RuntimeError(f"tensor {tensor} marked ready twice in one step (or out of range)")
        return [self._out[i] for i in range(n)]

    def pending(self) -> int:
        return self._lib.hvd_tracker_pending(self._h)

    def done(self, bucket: int):
        self._lib.hvd_tracker_bucket_done(self._h, bucket)

    def oldest_inflight_seconds(self) -> float:
        return self._lib.hvd_tracker_oldest_inflight_seconds(self._h)

    def close(self):
        if self._h:
            self._lib.hvd_tracker_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
