"""Chrome-trace timeline of the fused-gradient pipeline (the HOROVOD_TIMELINE role; Horovod's
timeline is external to the reference tree and not enabled by it — SURVEY.md §5.1). The writer is the C++
one in csrc/core/hvd_core.cpp; this is its Python handle. Open the JSON in chrome://tracing or Perfetto."""
from __future__ import annotations

import os

from .. import _native


class Timeline:
    def __init__(self, path: str, rank: int = 0):
        if rank != 0:
            root, ext = os.path.splitext(path)
            path = f"{root}.rank{rank}{ext}"
        self._lib = _native.host()
        self._h = self._lib.hvd_timeline_open(path.encode(), int(rank))
        if not self._h:
            raise OSError(f"cannot open timeline file {path}")
        self.path = path

    def now_us(self) -> float:
        return self._lib.hvd_timeline_now_us(self._h)

    def event(self, name: str, phase: str, lane: int, start_us: float, dur_us: float):
        self._lib.hvd_timeline_event(self._h, name.encode(), phase.encode(), int(lane), float(start_us), float(dur_us))

    def mark(self, name: str, lane: int = 0):
        self._lib.hvd_timeline_mark(self._h, name.encode(), int(lane), self.now_us())

    def close(self):
        if self._h:
            self._lib.hvd_timeline_close(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def from_env(rank: int = 0):
    """HOROVOD_TIMELINE=<file> (Horovod's env name) or HVD_TIMELINE."""
    path = os.environ.get("HOROVOD_TIMELINE") or os.environ.get("HVD_TIMELINE")
    return Timeline(path, rank) if path else None
