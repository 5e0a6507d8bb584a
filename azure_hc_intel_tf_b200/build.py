"""In-tree build of the two native libraries.

* ``libhvd_core.so``  — host-only C++ runtime (fusion planner, tracker, timeline, stall inspector).
* ``libb200cnn.so``   — every sm_100a CUDA kernel plus the CUDA-side runtime (symmetric memory,
  tensor-map encoding), compiled with ``-gencode arch=compute_100a,code=sm_100a -lineinfo``.

Both expose a plain C ABI declared in ``include/*.h`` and are loaded with ctypes, so the build
has no dependency on torch headers and finishes in seconds per file. The libraries live under
``azure_hc_intel_tf_b200/_lib/`` (git-ignored, but shipped to the GPU box by gpurun).

Staleness is decided by a content hash of the sources, not by mtimes: a snapshot copy to another
machine changes mtimes but must not trigger a rebuild.

Role in the reference: the install layer (``install-scripts/setup.sh`` and friends, ~80 min per
README.md:23-24) — collapsed here into one function that takes about a minute.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG = Path(__file__).resolve().parent
ROOT = PKG.parent
CSRC = PKG / "csrc"
INCLUDE = ROOT / "include"
LIBDIR = PKG / "_lib"
OBJDIR = LIBDIR / "obj"

HOST_LIB = LIBDIR / "libhvd_core.so"
CUDA_LIB = LIBDIR / "libb200cnn.so"

NVCC_ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr",
              "-Xptxas", "-v", "-I", str(INCLUDE), "-I", str(CSRC)]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; the CUDA library cannot be built")


def _digest(paths) -> str:
    h = hashlib.sha256()
    for p in sorted(paths):
        h.update(str(p.relative_to(ROOT)).encode())
        h.update(p.read_bytes())
    return h.hexdigest()


def host_sources():
    return sorted((CSRC / "core").glob("*.cpp"))


def cuda_sources():
    out = []
    for sub in ("kernels", "comm", "conv"):
        out += sorted((CSRC / sub).glob("*.cu"))
        out += sorted((CSRC / sub).glob("*.cpp"))
    return out


def _headers():
    return sorted(INCLUDE.glob("*.h")) + sorted(CSRC.rglob("*.cuh")) + sorted(CSRC.rglob("*.h"))


def _is_fresh(lib: Path, digest: str) -> bool:
    stamp = lib.with_suffix(".sha256")
    return lib.exists() and stamp.exists() and stamp.read_text().strip() == digest


def _run(cmd, log: Path | None = None):
    proc = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if log is not None:
        log.write_text(" ".join(map(str, cmd)) + "\n" + proc.stdout)
    if proc.returncode != 0:
        sys.stderr.write(proc.stdout)
        raise RuntimeError(f"build step failed ({proc.returncode}): {' '.join(map(str, cmd))}")
    return proc.stdout


def build_host(force: bool = False, verbose: bool = False) -> Path:
    srcs = host_sources()
    digest = _digest(srcs + _headers_host())
    if not force and _is_fresh(HOST_LIB, digest):
        return HOST_LIB
    LIBDIR.mkdir(parents=True, exist_ok=True)
    cxx = os.environ.get("CXX", "g++")
    cmd = [cxx, "-O2", "-std=c++17", "-shared", "-fPIC", "-Wall", "-I", str(INCLUDE)] + \
          [str(s) for s in srcs] + ["-o", str(HOST_LIB), "-lpthread"]
    if verbose:
        print(" ".join(cmd))
    _run(cmd)
    HOST_LIB.with_suffix(".sha256").write_text(digest)
    return HOST_LIB


def _headers_host():
    return sorted(INCLUDE.glob("hvd_core.h"))


def build_cuda(force: bool = False, verbose: bool = False, jobs: int | None = None) -> Path:
    srcs = cuda_sources()
    if not srcs:
        raise RuntimeError("no CUDA sources found")
    digest = _digest(srcs + _headers())
    if not force and _is_fresh(CUDA_LIB, digest):
        return CUDA_LIB
    OBJDIR.mkdir(parents=True, exist_ok=True)
    nvcc = _nvcc()

    def compile_one(src: Path) -> Path:
        obj = OBJDIR / (src.parent.name + "_" + src.stem + ".o")
        one = _digest([src] + _headers())
        stamp = obj.with_suffix(".sha256")
        if not force and obj.exists() and stamp.exists() and stamp.read_text().strip() == one:
            return obj
        cmd = [nvcc] + NVCC_ARCH + NVCC_FLAGS + ["-c", str(src), "-o", str(obj)]
        if src.suffix == ".cpp":
            cmd = [nvcc, "-O2", "-std=c++17", "-Xcompiler", "-fPIC", "-I", str(INCLUDE), "-I", str(CSRC),
                   "-x", "cu"] + NVCC_ARCH + ["-c", str(src), "-o", str(obj)]
        if verbose:
            print(" ".join(cmd))
        _run(cmd, log=obj.with_suffix(".log"))
        stamp.write_text(one)
        return obj

    jobs = jobs or min(8, os.cpu_count() or 4)
    with ThreadPoolExecutor(max_workers=jobs) as ex:
        objs = list(ex.map(compile_one, srcs))
    cmd = [nvcc, "-shared"] + NVCC_ARCH + [str(o) for o in objs] + ["-o", str(CUDA_LIB), "-lcuda"]
    if verbose:
        print(" ".join(cmd))
    _run(cmd)
    CUDA_LIB.with_suffix(".sha256").write_text(digest)
    return CUDA_LIB


def ptxas_report() -> str:
    """Concatenated ``-Xptxas -v`` output of the last build (registers / spills / smem)."""
    chunks = []
    for log in sorted(OBJDIR.glob("*.log")):
        chunks.append(f"==== {log.name}\n{log.read_text()}")
    return "\n".join(chunks)


def build_all(force: bool = False, verbose: bool = False):
    return build_host(force, verbose), build_cuda(force, verbose)


if __name__ == "__main__":
    force = "--force" in sys.argv
    h, c = build_all(force=force, verbose=True)
    print("built", h)
    print("built", c)
