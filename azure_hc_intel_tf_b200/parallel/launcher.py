"""Single-box launcher: the ``run-tf-sing-*.sh`` + ``mpirun`` role.

Reference contract (benchmark-scripts/run-tf-sing-ucx-openmpi.sh):

* four positionals ``NUM_NODES WORKERS_PER_SOCKET BATCH_SIZE FABRIC`` (:27-30);
* constants ``NUM_WARMUP_BATCHES=50 NUM_BATCHES=100 MODEL=resnet50 INTER_T=2`` (:32-35) — here also
  overridable through the environment instead of editing the script;
* topology from ``lscpu`` (:37-38) and the worker arithmetic at :40-50;
* the banner at :52-58; the command echo at :111;
* ``mpirun --map-by ppr:W:socket,pe=C`` process placement (:99-102) with
  ``HOROVOD_FUSION_THRESHOLD=134217728`` exported to every rank (:105).

Here a "socket" is a NUMA node, a "worker" is one process driving one GPU, the fabric is NVLink
(``nvl``; ``ib`` is accepted as an alias), NCCL (baseline) or gloo/TCP (``sock``). Ranks are spawned
locally with RANK/LOCAL_RANK/WORLD_SIZE/MASTER_ADDR/MASTER_PORT set, pinned to the cores of their socket,
and the whole group is torn down if any rank exits non-zero (mpirun's behaviour).
"""
from __future__ import annotations

import ctypes as C
import os
import re
import signal
import socket
import subprocess
import sys
import time
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence

from .. import _native

FUSION_THRESHOLD = 134217728


@dataclass
class WorkerLayout:
    num_nodes: int
    workers_per_socket: int
    num_sockets: int
    cores_per_socket: int
    inter_threads: int
    workers_per_node: int
    cores_per_worker: int
    intra_threads: int
    total_workers: int

    def banner(self) -> List[str]:
        return [f"TOTAL_NODES: {self.num_nodes}",
                f"WORKERS_PER_NODE: {self.workers_per_node}",
                f"TOTAL_WORKERS: {self.total_workers}",
                f"CORES_PER_WORKER: {self.cores_per_worker}",
                f"OMP_NUM_THREADS: {self.intra_threads}",
                f"NUM_INTRA_THREADS: {self.intra_threads}",
                f"NUM_INTER_THREADS: {self.inter_threads}"]


def worker_layout(num_nodes: int, workers_per_socket: int, num_sockets: int, cores_per_socket: int,
                  inter_threads: int = 2) -> WorkerLayout:
    out = (C.c_int32 * 4)()
    rc = _native.host().hvd_launch_arith(num_nodes, workers_per_socket, num_sockets, cores_per_socket,
                                         inter_threads, out)
    if rc != 0:
        raise ValueError("invalid launcher arguments")
    return WorkerLayout(num_nodes, workers_per_socket, num_sockets, cores_per_socket, inter_threads,
                        out[0], out[1], out[2], out[3])


def parse_lscpu(text: str) -> Dict[str, int]:
    """``Socket(s)`` and ``Core(s) per socket`` as the reference greps them (:37-38)."""
    def grab(key):
        m = re.search(rf"^{re.escape(key)}\s*:\s*(\d+)", text, re.M)
        return int(m.group(1)) if m else None
    sockets = grab("Socket(s)") or 1
    cores = grab("Core(s) per socket")
    if cores is None:
        cores = max(1, (os.cpu_count() or 1) // sockets)
    return {"sockets": sockets, "cores_per_socket": cores}


def detect_cpu() -> Dict[str, int]:
    try:
        txt = subprocess.run(["lscpu"], capture_output=True, text=True, timeout=10).stdout
    except Exception:
        txt = ""
    return parse_lscpu(txt)


def socket_cpus(sock: int, num_sockets: int) -> List[int]:
    """CPU ids of a socket: NUMA node cpulist when sysfs has it, else an even split of the affinity mask."""
    path = f"/sys/devices/system/node/node{sock}/cpulist"
    avail = sorted(os.sched_getaffinity(0))
    if num_sockets > 1 and os.path.exists(path):
        cpus = []
        for part in open(path).read().strip().split(","):
            if "-" in part:
                a, b = part.split("-")
                cpus += list(range(int(a), int(b) + 1))
            elif part:
                cpus.append(int(part))
        cpus = [c for c in cpus if c in avail]
        if cpus:
            return cpus
    per = max(1, len(avail) // num_sockets)
    return avail[sock * per:(sock + 1) * per] or avail


def free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def benchmark_args(batch_size: int, model: str, num_batches: int, num_warmup: int, layout: WorkerLayout,
                   fabric: str, device: str, extra: Sequence[str] = ()) -> List[str]:
    """The TF_ARGS block of the reference (:62-81), flag for flag."""
    args = [f"--batch_size={batch_size}", f"--num_warmup_batches={num_warmup}", f"--num_batches={num_batches}",
            f"--model={model}", f"--num_intra_threads={layout.intra_threads}",
            f"--num_inter_threads={layout.inter_threads}", "--kmp_blocktime=1",
            "--kmp_affinity=granularity=fine,noverbose,compact,1,0", "--display_every=10", "--data_format=NCHW",
            "--optimizer=momentum", "--forward_only=False", f"--device={device}", "--mkl=TRUE",
            "--variable_update=horovod", f"--horovod_device={device}", f"--local_parameter_device={device}",
            "--data_name=imagenet", f"--fabric={fabric}"]
    return args + list(extra)


def launch(num_nodes: int, workers_per_socket: int, batch_size: int, fabric: str, *, model: str = "resnet50",
           num_batches: int = 100, num_warmup: int = 50, inter_threads: int = 2, device: Optional[str] = None,
           extra: Sequence[str] = (), node_rank: int = 0, master_addr: str = "127.0.0.1",
           master_port: Optional[int] = None, dry_run: bool = False, module: str = "azure_hc_intel_tf_b200.cli.cnn_benchmarks",
           max_workers: Optional[int] = None, out=sys.stdout) -> int:
    cpu = detect_cpu()
    layout = worker_layout(num_nodes, workers_per_socket, cpu["sockets"], cpu["cores_per_socket"], inter_threads)
    if device is None:
        try:
            import torch
            device = "gpu" if torch.cuda.is_available() and fabric not in ("sock", "gloo") else "cpu"
        except Exception:
            device = "cpu"
    wpn = layout.workers_per_node
    if device == "gpu":
        import torch
        ngpu = torch.cuda.device_count()
        if wpn > ngpu:
            print(f"WORKERS_PER_NODE {wpn} exceeds the {ngpu} visible GPUs; clamping", file=out)
            wpn = ngpu
    if max_workers:
        wpn = min(wpn, max_workers)
    total = num_nodes * wpn
    for line in layout.banner():
        print(line, file=out)
    topo = None
    if device == "gpu":
        # fabric health probe (the reference runs `ibv_devinfo | grep state` on every node: prep-cluster.sh:23)
        try:
            from . import topology
            topo = topology.discover()
            for line in topo.report():
                print(line, file=out)
            if fabric in ("nvl", "ib") and len(topo.gpus) > 1 and not topo.all_to_all_p2p:
                print("fabric nvl needs all-to-all peer access; falling back to nccl", file=out)
                fabric = "nccl"
        except Exception as e:   # discovery is advisory
            print(f"topology discovery skipped: {e}", file=out)
    master_port = master_port or int(os.environ.get("MASTER_PORT", 0)) or free_port()
    args = benchmark_args(batch_size, model, num_batches, num_warmup, layout, fabric, device, extra)
    cmd = [sys.executable, "-m", module] + args
    print("Common Args: " + " ".join(args), file=out)
    print(f"Fabric: {fabric}", file=out)
    print(" ".join(cmd), file=out, flush=True)
    if dry_run:
        return 0
    procs: List[subprocess.Popen] = []
    per_socket = max(1, wpn // cpu["sockets"]) if workers_per_socket else 1
    for lr in range(wpn):
        env = dict(os.environ)
        env.update(RANK=str(node_rank * wpn + lr), LOCAL_RANK=str(lr), WORLD_SIZE=str(total),
                   LOCAL_WORLD_SIZE=str(wpn), MASTER_ADDR=master_addr, MASTER_PORT=str(master_port),
                   OMP_NUM_THREADS=str(max(1, layout.intra_threads)))
        # The reference exports HOROVOD_FUSION_THRESHOLD=134217728 to every rank (:105). A value the user set is
        # forwarded as is; otherwise the library fabrics get the reference's 128 MiB and the NVLink fabric keeps its
        # own default (16 MiB buckets that overlap with backward; utils/flags.py resolve_fusion_threshold).
        if "HOROVOD_FUSION_THRESHOLD" not in os.environ and fabric not in ("nvl", "ib"):
            env["HOROVOD_FUSION_THRESHOLD"] = str(FUSION_THRESHOLD)
        sock = min(lr // per_socket, cpu["sockets"] - 1) if workers_per_socket else 0
        if topo is not None and lr < len(topo.gpus) and 0 <= topo.numa_of(lr) < cpu["sockets"]:
            sock = topo.numa_of(lr)      # pin the worker next to the NUMA node its GPU hangs off
        cpus = socket_cpus(sock, cpu["sockets"]) if workers_per_socket else sorted(os.sched_getaffinity(0))
        # the pe=<cores> part of --map-by: give each worker its own slice of the socket's cores
        k = lr % per_socket
        width = max(1, len(cpus) // per_socket)
        mine = cpus[k * width:(k + 1) * width] or cpus

        def pin(mask=tuple(mine)):
            try:
                os.sched_setaffinity(0, mask)
            except OSError:
                pass
            os.setpgrp()

        procs.append(subprocess.Popen(cmd, env=env, preexec_fn=pin))
    return _wait_group(procs)


def _wait_group(procs: List[subprocess.Popen]) -> int:
    """mpirun semantics: the first non-zero exit kills every other rank."""
    rc = 0
    alive = list(procs)
    try:
        while alive:
            for p in list(alive):
                r = p.poll()
                if r is None:
                    continue
                alive.remove(p)
                if r != 0 and rc == 0:
                    rc = r
                    for q in alive:
                        try:
                            os.killpg(q.pid, signal.SIGTERM)
                        except ProcessLookupError:
                            pass
            time.sleep(0.05)
    except KeyboardInterrupt:
        for q in alive:
            try:
                os.killpg(q.pid, signal.SIGTERM)
            except ProcessLookupError:
                pass
        rc = 130
    return rc


def main(argv=None) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    if len(argv) < 4 or argv[0] in ("-h", "--help"):
        print("usage: launcher <NUM_NODES> <WORKERS_PER_SOCKET> <batch_size> <fabric(nvl|ib|nccl|sock)> [extra flags]\n"
              "  If WORKERS_PER_SOCKET=0, then WORKERS_PER_NODE=1\n"
              "  env overrides: MODEL NUM_BATCHES NUM_WARMUP_BATCHES INTER_T NODE_RANK MASTER_ADDR MASTER_PORT DRY_RUN")
        return 2
    nn, wps, bs, fabric = int(argv[0]), int(argv[1]), int(argv[2]), argv[3]
    env = os.environ
    return launch(nn, wps, bs, fabric, model=env.get("MODEL", "resnet50"),
                  num_batches=int(env.get("NUM_BATCHES", 100)), num_warmup=int(env.get("NUM_WARMUP_BATCHES", 50)),
                  inter_threads=int(env.get("INTER_T", 2)), extra=argv[4:], node_rank=int(env.get("NODE_RANK", 0)),
                  master_addr=env.get("MASTER_ADDR", "127.0.0.1"), dry_run=env.get("DRY_RUN", "") not in ("", "0"),
                  max_workers=int(env["MAX_WORKERS"]) if env.get("MAX_WORKERS") else None)


if __name__ == "__main__":
    sys.exit(main())
