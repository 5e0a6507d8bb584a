"""OSU-style latency / bandwidth sweeps over the NVLink fabric (the osu_allreduce, osu_allgather, osu_bcast,
osu_barrier, osu_latency, osu_bw and osu_bibw roles).

The reference installs OSU micro-benchmarks 5.6.1 (install-scripts/install_osu_bench.sh:13-17; image definition
tf-hvd-gcc-ompi-ucx-mlnx-osu.def:25-26,58-62) but never runs them; BASELINE.json turns that role into a built-in
sweep: 1 KB – 1 GB at 2/4/8 GPUs, this framework's NVLink kernels next to NCCL (``torch.distributed``), timed on the
device (CUDA events, or %globaltimer inside the ping-pong kernel), MAX over ranks.

    torchrun --nproc-per-node N -m azure_hc_intel_tf_b200.bench.allreduce_sweep --op allreduce [--algos auto,oneshot,twoshot,nvls]
    torchrun --nproc-per-node N -m azure_hc_intel_tf_b200.bench.allreduce_sweep --op allgather|broadcast|barrier
    torchrun --nproc-per-node N -m azure_hc_intel_tf_b200.bench.allreduce_sweep --op p2p        # latency, bw, bibw

Every point is the MEDIAN of ``--trials`` timed batches (a single batch is at the mercy of one slow launch; round 1's
"4 MiB cliff" was such an outlier). Columns: bytes, µs (ours per algorithm / NCCL), bus bandwidth GB/s
(allreduce: algbw·2(N−1)/N; allgather: algbw·(N−1)/N with algbw = N·segment / time; broadcast, p2p: bytes / time).
``--op allreduce`` with several ``--algos`` also prints the size thresholds at which each algorithm starts to win,
which is where parallel/symm.py's defaults (B200_COMM_ONE_SHOT_MAX, B200_COMM_NVLS_MIN) come from.
"""
from __future__ import annotations

import argparse
import json
import sys

import torch
import torch.distributed as dist


def timed(fn, iters, warmup, trials=5):
    """Median over ``trials`` of (device time of ``iters`` back-to-back calls / iters), MAX over ranks, in µs."""
    for _ in range(warmup):
        fn()
    out = []
    for _ in range(trials):
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / iters * 1e3], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out.append(float(t.item()))
    out.sort()
    return out[len(out) // 2]


def iters_for(nbytes):
    return 200 if nbytes <= (1 << 20) else (40 if nbytes <= (1 << 26) else 8)


def sizes(lo, hi, step=4):
    s = lo
    while s <= hi:
        yield s
        s *= step


def sweep_allreduce(ctx, a, rank, n_ranks, rows):
    from ..parallel import symm as symm_mod
    algos = [x for x in a.algos.split(",") if x]
    f = 2.0 * (n_ranks - 1) / n_ranks
    if rank == 0:
        print(f"# allreduce sweep, {n_ranks} ranks, device-timed, median of {a.trials} batches, max over ranks; auto = one-shot "
              f"<= {ctx.one_shot_max_bytes} B, NVLS >= {ctx.nvls_min_bytes} B (fp32 with a multicast mapping), else two-shot")
        print(f"{'dtype':9s} {'bytes':>11s} " + " ".join(f"{x + ' us':>11s}" for x in algos) + f" {'nccl us':>9s} "
              f"{'best busbw':>10s} {'nccl busbw':>10s}")
    for dname in a.dtypes.split(","):
        dtype = getattr(torch, dname)
        esz = torch.empty((), dtype=dtype).element_size()
        gran = 1024 if dtype == torch.float32 else 2048
        buf = ctx.alloc(a.max + gran * esz, dtype, multicast=(dtype == torch.float32))
        has_mc = bool(getattr(buf, "mc_ptr", 0))
        for size in sizes(a.min, a.max):
            n = size // esz
            small = dtype == torch.float32 and n * esz <= 256 * 1024        # one-shot range: any multiple of 4 elements
            if not small:
                n = max(gran, n // gran * gran)
            nbytes = n * esz
            it = iters_for(nbytes)
            row = {"op": "allreduce", "dtype": dname, "bytes": nbytes}
            for algo in algos:
                if dtype != torch.float32 and algo != "auto":
                    continue
                if algo == "nvls" and (not has_mc or nbytes < 64 * 1024):
                    continue
                if algo == "oneshot" and nbytes > 256 * 1024:
                    continue
                if algo == "twoshot" and n % gran:
                    continue
                code = symm_mod.ALGOS[algo]
                if algo == "auto" and n % gran and nbytes > ctx.one_shot_max_bytes:
                    continue
                row[algo + "_us"] = timed(lambda: ctx.allreduce_range_(buf, 0, n, 1.0 / n_ranks, algo=code), it, 5, a.trials)
            x = torch.ones(n, device="cuda", dtype=dtype)
            row["nccl_us"] = timed(lambda: dist.all_reduce(x), it, 5, a.trials)
            del x
            ours = [row[k] for k in row if k.endswith("_us") and k != "nccl_us"]
            row["ours_us"] = row.get("auto_us", min(ours))
            row["ours_busbw_gbs"] = nbytes / row["ours_us"] / 1e3 * f
            row["best_busbw_gbs"] = nbytes / min(ours) / 1e3 * f
            row["nccl_busbw_gbs"] = nbytes / row["nccl_us"] / 1e3 * f
            rows.append(row)
            if rank == 0:
                cells = " ".join(f"{row[x + '_us']:11.1f}" if x + "_us" in row else f"{'--':>11s}" for x in algos)
                print(f"{dname:9s} {nbytes:11d} {cells} {row['nccl_us']:9.1f} {row['best_busbw_gbs']:10.1f} "
                      f"{row['nccl_busbw_gbs']:10.1f}", flush=True)
        del buf
    # thresholds from the data: the smallest size from which an algorithm is never again slower than its neighbour
    fr = [r for r in rows if r["dtype"] == "float32"]
    th = {}
    if "oneshot" in algos and "twoshot" in algos:
        both = [r for r in fr if "oneshot_us" in r and "twoshot_us" in r]
        win = [r["bytes"] for r in both if r["oneshot_us"] <= r["twoshot_us"]]
        th["one_shot_max_bytes"] = max(win) if win else 0
    both = [r for r in fr if "nvls_us" in r and "twoshot_us" in r]
    if both:
        never = [r["bytes"] for r in both if r["nvls_us"] > r["twoshot_us"]]
        ok = [r["bytes"] for r in both if r["bytes"] > (max(never) if never else 0)]
        th["nvls_min_bytes"] = min(ok) if ok else -1
    if rank == 0 and th:
        print(f"# thresholds derived from this sweep: {th}")
    return th


def sweep_simple(ctx, a, rank, n_ranks, rows):
    """allgather / broadcast / barrier."""
    if a.op == "barrier":
        ours = timed(ctx.barrier, 500, 20, a.trials)
        one = torch.zeros(1, device="cuda")
        nccl = timed(lambda: dist.all_reduce(one), 500, 20, a.trials)     # the smallest NCCL collective = its barrier cost
        rows.append({"op": "barrier", "ours_us": ours, "nccl_us": nccl})
        if rank == 0:
            print(f"# barrier, {n_ranks} ranks, device-timed, max over ranks")
            print(f"flag barrier {ours:.1f} us    NCCL 4-byte allreduce {nccl:.1f} us")
        return
    if rank == 0:
        what = ("root 0 writes every peer's buffer" if a.op == "broadcast" else
                "every rank writes its segment into every peer's buffer; bytes = total gathered")
        print(f"# {a.op} sweep ({what}; multicast stores when the buffer has an NVLS mapping), {n_ranks} ranks, "
              f"device-timed, median of {a.trials}, max over ranks")
        print(f"{'bytes':>12s} {'ours us':>10s} {'nccl us':>10s} {'ours GB/s':>11s} {'nccl GB/s':>11s}")
    buf = ctx.alloc(a.max + 4096, torch.float32, multicast=True)
    for size in sizes(a.min, a.max):
        if a.op == "broadcast":
            n = max(4, size // 4 // 4 * 4)
            nbytes = n * 4
            x = torch.ones(n, device="cuda")
            ours = timed(lambda: ctx.broadcast_(buf, 0, n, 0), iters_for(nbytes), 5, a.trials)
            nccl = timed(lambda: dist.broadcast(x, src=0), iters_for(nbytes), 5, a.trials)
            bw = 1.0
        else:
            seg = max(4, size // 4 // n_ranks // 4 * 4)
            nbytes = seg * 4 * n_ranks
            x = torch.ones(seg, device="cuda")
            out = torch.empty(seg * n_ranks, device="cuda")
            ours = timed(lambda: ctx.allgather_(buf, 0, seg), iters_for(nbytes), 5, a.trials)
            nccl = timed(lambda: dist.all_gather_into_tensor(out, x), iters_for(nbytes), 5, a.trials)
            bw = (n_ranks - 1) / n_ranks
            del out
        row = {"op": a.op, "dtype": "float32", "bytes": nbytes, "ours_us": ours, "nccl_us": nccl,
               "ours_busbw_gbs": nbytes / ours / 1e3 * bw, "nccl_busbw_gbs": nbytes / nccl / 1e3 * bw}
        rows.append(row)
        if rank == 0:
            print(f"{nbytes:12d} {ours:10.1f} {nccl:10.1f} {row['ours_busbw_gbs']:11.1f} {row['nccl_busbw_gbs']:11.1f}", flush=True)
        del x


def sweep_p2p(ctx, a, rank, n_ranks, rows):
    """osu_latency / osu_bw / osu_bibw between GPU 0 and each other GPU (the fabric is an NVSwitch: every pair should
    read the same; BASELINE.md's 900 GB/s per direction is the nominal figure this confirms or corrects)."""
    buf = ctx.alloc(max(a.max, 4 << 20) + 4096, torch.float32)
    out = torch.zeros(2, device="cuda")
    peers = list(range(1, n_ranks)) if a.all_pairs else [1, n_ranks - 1] if n_ranks > 2 else [1]
    if rank == 0:
        print(f"# point-to-point over NVLink, {n_ranks} ranks; latency = ping-pong kernel, one-way, %globaltimer on the "
              f"device; bandwidth = peer-store / peer-load kernel, CUDA events, median of {a.trials}")
        print(f"{'pair':>7s} {'bytes':>11s} {'lat us':>8s} {'push GB/s':>10s} {'pull GB/s':>10s} {'bidir GB/s':>11s}")
    for peer in peers:
        for size in sizes(max(a.min, 16), a.max):
            n = max(4, size // 4 // 4 * 4)
            nbytes = n * 4
            row = {"op": "p2p", "pair": [0, peer], "bytes": nbytes}
            if nbytes <= (4 << 20):
                iters = 200 if nbytes <= 65536 else 30
                lat = []
                for _ in range(a.trials):
                    ctx.p2p_pingpong(buf, 0, peer, n, iters, out)
                    torch.cuda.synchronize()
                    t = out[:1].double().clone()
                    dist.all_reduce(t, op=dist.ReduceOp.MAX)      # only rank 0 measured; the others hold zero
                    lat.append(float(t.item()))
                lat.sort()
                row["latency_us"] = lat[len(lat) // 2]
            if nbytes >= 4096:
                it = iters_for(nbytes)

                def one(pull, both):
                    def fn():
                        if rank == 0:
                            ctx.p2p_copy(buf, peer, 0, n, pull=pull)
                        elif both and rank == peer:
                            ctx.p2p_copy(buf, 0, 0, n, pull=pull)
                    return timed(fn, it, 3, a.trials)
                row["push_gbs"] = nbytes / one(False, False) / 1e3
                row["pull_gbs"] = nbytes / one(True, False) / 1e3
                row["bidir_gbs"] = 2 * nbytes / one(False, True) / 1e3
            rows.append(row)
            if rank == 0:
                g = lambda k: f"{row[k]:.1f}" if k in row else "--"
                print(f"{'0-' + str(peer):>7s} {nbytes:11d} {g('latency_us'):>8s} {g('push_gbs'):>10s} {g('pull_gbs'):>10s} "
                      f"{g('bidir_gbs'):>11s}", flush=True)


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--min-bytes", dest="min", type=int, default=1024)
    ap.add_argument("--max-bytes", dest="max", type=int, default=1 << 30)
    ap.add_argument("--dtypes", default="float32,bfloat16")
    ap.add_argument("--algos", default="auto", help="allreduce: comma list of auto,oneshot,twoshot,nvls")
    ap.add_argument("--trials", type=int, default=5)
    ap.add_argument("--all-pairs", action="store_true", help="p2p: GPU 0 against every other GPU (default: first and last)")
    ap.add_argument("--json", default=None)
    ap.add_argument("--op", default="allreduce", choices=["allreduce", "allgather", "broadcast", "barrier", "p2p"])
    a = ap.parse_args(argv)
    from .. import hvd
    hvd.init(fabric="nvl")
    ctx = hvd.symmetric_context()
    n_ranks, rank = hvd.size(), hvd.rank()
    if ctx is None:
        raise SystemExit("the sweep needs at least 2 ranks on the nvl fabric")
    rows, extra = [], {}
    # everything is issued ON the context's communication stream: a collective called from another stream forks onto it
    # and joins back (two cross-stream event hops, ~15 us), which is the caller's cost, not the collective's
    with torch.cuda.stream(ctx.stream):
        if a.op == "allreduce":
            extra["thresholds"] = sweep_allreduce(ctx, a, rank, n_ranks, rows)
        elif a.op == "p2p":
            sweep_p2p(ctx, a, rank, n_ranks, rows)
        else:
            sweep_simple(ctx, a, rank, n_ranks, rows)
    if a.json and rank == 0:
        with open(a.json, "w") as f:
            json.dump({"op": a.op, "ranks": n_ranks, "trials": a.trials, "ctas": ctx.ctas, **extra, "rows": rows}, f, indent=1)
    hvd.shutdown()
    return 0


if __name__ == "__main__":
    sys.exit(main())
