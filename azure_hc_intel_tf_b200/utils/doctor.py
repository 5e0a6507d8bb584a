"""Environment report and sanity check: ``python -m azure_hc_intel_tf_b200.utils.doctor``.

Role in the reference stack: the ``%runscript`` of the Singularity images, which `build-container.sh` runs right after
the build as its sanity check (install-scripts/build-container.sh:29-30) and which prints the OS, gcc, TensorFlow
version, whether MKL is enabled, the Horovod version, the MPI and the OFED / libfabric versions
(install-scripts/tf-hvd-gcc-ompi-ucx-mlnx.def:45-55, tf-hvd-gcc-impi-libfabric-mlnx.def:41-53); plus the limits
`update_config.sh` sets (memlock unlimited, nofile 65535; install-scripts/update_config.sh:6-11) — checked here, not
changed. The same questions for this stack: which compiler / CUDA toolkit / driver / PyTorch / NCCL, do the two native
libraries build and load, which GPUs and is the NVLink fabric whole, are the process limits sane for pinned host
buffers and many ranks.

Exit code 0 = ready for the chosen mode, 1 = something required is missing (each problem is printed as ``FAIL``).
With no GPU the report says so and checks the CPU plumbing path only (the reference's ``sock`` fabric).
"""
from __future__ import annotations

import os
import platform
import resource
import shutil
import subprocess
import sys
from typing import List, Tuple


def _run(cmd: List[str]) -> str:
    try:
        return subprocess.run(cmd, capture_output=True, text=True, timeout=20).stdout.strip()
    except Exception:
        return ""


def _first_line(s: str) -> str:
    return s.splitlines()[0] if s else "not found"


def collect() -> Tuple[List[str], List[str]]:
    """Returns (report lines, failures)."""
    out: List[str] = []
    fail: List[str] = []

    def item(key: str, value: str):
        out.append(f"{key:<28s} {value}")

    item("OS", f"{platform.system()} {platform.release()} ({platform.machine()})")
    try:
        with open("/etc/os-release") as f:
            pretty = [l.split("=", 1)[1].strip().strip('"') for l in f if l.startswith("PRETTY_NAME=")]
        if pretty:
            item("distribution", pretty[0])
    except OSError:
        pass
    item("python", sys.version.split()[0])
    gxx = shutil.which("g++")
    item("g++", _first_line(_run([gxx, "--version"])) if gxx else "not found")
    if not gxx:
        fail.append("g++ is needed to build libhvd_core.so")
    nvcc = shutil.which("nvcc") or ("/usr/local/cuda/bin/nvcc" if os.path.exists("/usr/local/cuda/bin/nvcc") else None)
    if nvcc:
        rel = [l for l in _run([nvcc, "--version"]).splitlines() if "release" in l]
        item("nvcc", rel[0].strip() if rel else "found")
    else:
        item("nvcc", "not found")

    # framework
    try:
        import torch
        import torch.distributed as dist
        item("torch", f"{torch.__version__} (CUDA runtime {torch.version.cuda})")
        item("torch.distributed backends", ", ".join(b for b, ok in (("nccl", dist.is_nccl_available()),
                                                                      ("gloo", dist.is_gloo_available())) if ok))
        if torch.cuda.is_available():
            try:
                item("NCCL (baseline fabric)", ".".join(str(v) for v in torch.cuda.nccl.version()))
            except Exception:
                item("NCCL (baseline fabric)", "unknown")
        if not dist.is_gloo_available():
            fail.append("torch.distributed has no gloo backend (CPU plumbing path)")
    except Exception as e:                                   # pragma: no cover - torch is a hard dependency
        torch = None
        item("torch", f"import failed: {e}")
        fail.append("PyTorch is required")

    # native libraries
    try:
        from .. import _native, build
        host = _native.host()
        item("libhvd_core.so", host.hvd_core_version().decode())
    except Exception as e:
        item("libhvd_core.so", f"FAILED: {e}")
        fail.append("host runtime library does not build / load")
    have_gpu = bool(torch is not None and torch.cuda.is_available())
    try:
        from .. import _native, build
        if nvcc or build.CUDA_LIB.exists():
            lib = _native.cuda()
            ver = lib.b200_version().decode() if hasattr(lib, "b200_version") else "loaded"
            item("libb200cnn.so", f"{ver}; smallest Cin on the TMA path {lib.b200_conv_min_channels()}")
        else:
            item("libb200cnn.so", "not built (no nvcc here)")
            if have_gpu:
                fail.append("a GPU is present but the sm_100a kernel library is missing and nvcc is not available")
    except Exception as e:
        item("libb200cnn.so", f"FAILED: {e}")
        if have_gpu or nvcc:
            fail.append("kernel library does not build / load")

    # devices and fabric
    from ..parallel import topology
    topo = topology.discover()
    out.extend(topo.report())
    if have_gpu:
        for g in topo.gpus:
            if "B200" not in g.name and "B100" not in g.name and "GB200" not in g.name:
                fail.append(f"gpu{g.index} is '{g.name}': the kernels are built for sm_100a only")
        if len(topo.gpus) > 1 and not topo.all_to_all_p2p:
            fail.append("no all-to-all peer access: fabric nvl cannot be used (nccl still can)")
        if not topo.nvlink_healthy:
            fail.append("some NVLink links are down")
    else:
        item("mode", "no CUDA device visible: CPU plumbing only (fabric sock/gloo, --kernels=cudnn)")

    # process limits (update_config.sh role: checked, never changed)
    def lim(name, res, want):
        soft, hard = resource.getrlimit(res)
        s = "unlimited" if soft == resource.RLIM_INFINITY else str(soft)
        ok = soft == resource.RLIM_INFINITY or soft >= want
        item(f"ulimit {name}", f"{s}{'' if ok else f'   (recommended >= {want})'}")
        return ok

    lim("memlock (pinned batches)", resource.RLIMIT_MEMLOCK, 1 << 30)
    if not lim("nofile", resource.RLIMIT_NOFILE, 4096):
        out.append("  note: a low nofile limit can break 8-rank rendezvous + IPC handle exchange")
    item("HOROVOD_FUSION_THRESHOLD", os.environ.get("HOROVOD_FUSION_THRESHOLD", "unset (default 134217728)"))
    return out, fail


def main() -> int:
    lines, fail = collect()
    print("B200 data-parallel CNN benchmark: environment")
    print("\n".join(lines))
    for f in fail:
        print(f"FAIL  {f}")
    print("status:", "ready" if not fail else f"{len(fail)} problem(s)")
    return 0 if not fail else 1


if __name__ == "__main__":
    sys.exit(main())
