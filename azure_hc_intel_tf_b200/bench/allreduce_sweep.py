"""OSU-style collective latency / bandwidth sweep (the osu_allreduce, osu_bcast and osu_barrier roles).

The reference installs OSU micro-benchmarks 5.6.1 (install-scripts/install_osu_bench.sh:13-17) but never runs them;
BASELINE.json turns that role into a built-in sweep: 1 KB – 1 GB, fp32 and bf16, this framework's NVLink kernels
next to NCCL (``torch.distributed.all_reduce``), device-timed, MAX over ranks.

    torchrun --nproc-per-node N -m azure_hc_intel_tf_b200.bench.allreduce_sweep [--min-bytes 1024] [--max-bytes 1073741824]

Columns: bytes, latency µs (ours / NCCL), algbw GB/s = bytes/time, busbw = algbw·2(N−1)/N.
``--op broadcast`` sweeps the root-writes-to-all broadcast kernel (fp32) against ``dist.broadcast`` (busbw = algbw),
``--op barrier`` times the flag barrier against ``dist.barrier``'s NCCL implementation (one row, no payload).
"""
from __future__ import annotations

import argparse
import json
import sys

import torch
import torch.distributed as dist


def timed(fn, iters, warmup):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters * 1e3], device="cuda", dtype=torch.float64)   # µs
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--min-bytes", dest="min", type=int, default=1024)
    ap.add_argument("--max-bytes", dest="max", type=int, default=1 << 30)
    ap.add_argument("--dtypes", default="float32,bfloat16")
    ap.add_argument("--json", default=None)
    ap.add_argument("--op", default="allreduce", choices=["allreduce", "broadcast", "barrier"])
    a = ap.parse_args(argv)
    from .. import hvd
    hvd.init(fabric="nvl")
    ctx = hvd.symmetric_context()
    n_ranks, rank = hvd.size(), hvd.rank()
    if ctx is None:
        raise SystemExit("the sweep needs at least 2 ranks on the nvl fabric")
    rows = []
    if a.op == "barrier":
        ours = timed(ctx.barrier, 500, 20)
        one = torch.zeros(1, device="cuda")
        nccl = timed(lambda: dist.all_reduce(one), 500, 20)      # the smallest NCCL collective = its barrier cost
        rows.append({"op": "barrier", "ours_us": ours, "nccl_us": nccl})
        if rank == 0:
            print(f"# barrier, {n_ranks} ranks, device-timed, max over ranks")
            print(f"flag barrier {ours:.1f} us    NCCL 4-byte allreduce {nccl:.1f} us")
    if a.op == "broadcast":
        if rank == 0:
            print(f"# broadcast sweep (root 0 writes every peer's buffer over NVLink), {n_ranks} ranks, device-timed, max over ranks")
            print(f"{'dtype':9s} {'bytes':>12s} {'ours us':>10s} {'nccl us':>10s} {'ours GB/s':>11s} {'nccl GB/s':>11s}")
        buf = ctx.alloc(a.max + 4096, torch.float32)
        size = a.min
        while size <= a.max:
            n = max(1024, size // 4 // 1024 * 1024)
            nbytes = n * 4
            iters = 200 if nbytes <= (1 << 20) else (40 if nbytes <= (1 << 26) else 10)
            x = torch.ones(n, device="cuda")
            ours = timed(lambda: ctx.broadcast_(buf, 0, n, 0), iters, 5)
            nccl = timed(lambda: dist.broadcast(x, src=0), iters, 5)
            row = {"op": "broadcast", "dtype": "float32", "bytes": nbytes, "ours_us": ours, "nccl_us": nccl,
                   "ours_busbw_gbs": nbytes / ours / 1e3, "nccl_busbw_gbs": nbytes / nccl / 1e3}
            rows.append(row)
            if rank == 0:
                print(f"{'float32':9s} {nbytes:12d} {ours:10.1f} {nccl:10.1f} {row['ours_busbw_gbs']:11.1f} {row['nccl_busbw_gbs']:11.1f}",
                      flush=True)
            size *= 4
            del x
    if a.op != "allreduce":
        if a.json and rank == 0:
            with open(a.json, "w") as f:
                json.dump({"ranks": n_ranks, "rows": rows}, f, indent=1)
        hvd.shutdown()
        return 0
    if rank == 0:
        print(f"# allreduce sweep, {n_ranks} ranks, device-timed, max over ranks; one-shot <= {ctx.one_shot_max_bytes} B, "
              f"NVLS >= {ctx.nvls_min_bytes} B (fp32, when a multicast mapping exists)")
        print(f"{'dtype':9s} {'bytes':>12s} {'ours us':>10s} {'nccl us':>10s} {'ours busbw':>11s} {'nccl busbw':>11s}")
    for dname in a.dtypes.split(","):
        dtype = getattr(torch, dname)
        esz = torch.empty((), dtype=dtype).element_size()
        gran = 1024 if dtype == torch.float32 else 2048
        buf = ctx.alloc(a.max + gran * esz, dtype, multicast=(dtype == torch.float32))
        size = a.min
        while size <= a.max:
            n = max(gran, size // esz // gran * gran)
            nbytes = n * esz
            iters = 200 if nbytes <= (1 << 20) else (40 if nbytes <= (1 << 26) else 10)
            x = torch.ones(n, device="cuda", dtype=dtype)
            ours = timed(lambda: ctx.allreduce_range_(buf, 0, n, 1.0 / n_ranks), iters, 5)
            nccl = timed(lambda: dist.all_reduce(x), iters, 5)
            f = 2.0 * (n_ranks - 1) / n_ranks
            row = {"dtype": dname, "bytes": nbytes, "ours_us": ours, "nccl_us": nccl,
                   "ours_busbw_gbs": nbytes / ours / 1e3 * f, "nccl_busbw_gbs": nbytes / nccl / 1e3 * f}
            rows.append(row)
            if rank == 0:
                print(f"{dname:9s} {nbytes:12d} {ours:10.1f} {nccl:10.1f} {row['ours_busbw_gbs']:11.1f} {row['nccl_busbw_gbs']:11.1f}",
                      flush=True)
            size *= 4
        del x
    if a.json and rank == 0:
        with open(a.json, "w") as f:
            json.dump({"ranks": n_ranks, "rows": rows}, f, indent=1)
    hvd.shutdown()
    return 0


if __name__ == "__main__":
    sys.exit(main())
