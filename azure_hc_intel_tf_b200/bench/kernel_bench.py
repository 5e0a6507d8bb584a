"""Per-layer kernel micro-benchmark: every distinct convolution geometry of a model through the tcgen05
kernels and through the library (cuDNN via aten), forward / backward-data / backward-weights, plus the BN kernels.

    python -m azure_hc_intel_tf_b200.bench.kernel_bench --model resnet50 --batch 256 [--json out.json]

Timing: CUDA events around ``iters`` back-to-back launches after warm-up; a 256 MiB buffer is rewritten before
each timed group so operands do not start in L2. Roofline columns use MEASURED_PEAKS.json when present
(bf16_tflops_sustained, hbm_gbs), else the profiling guide's fallback (1590 TFLOP/s, 6650 GB/s).

Reference role: the per-primitive view of MKL-DNN verbose mode for the reference's CPU stack; used here to decide
which kernel to optimise next and to fill profiles/.
"""
from __future__ import annotations

import argparse
import json
import os
from collections import OrderedDict

import torch

from .. import models
from ..models.builder import Conv
from ..ops import b200, conv_impl, conv_umma


def peaks():
    root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    try:
        with open(os.path.join(root, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return p.get("bf16_tflops_sustained", p["bf16_tflops"]), p["hbm_gbs"], "measured"
    except Exception:
        return 1590.0, 6650.0, "fallback"


_flush = None


def flush_l2():
    global _flush
    if _flush is None:
        _flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    _flush.zero_()


def timeit(fn, iters=5, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush_l2()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]   # median, microseconds


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="resnet50")
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--json", default=None)
    ap.add_argument("--no-lib", action="store_true")
    ap.add_argument("--no-cudnn-benchmark", action="store_true",
                    help="leave cuDNN on its heuristics (default: benchmark mode, i.e. per-shape algorithm search - the "
                         "library column is then the library at its best)")
    a = ap.parse_args(argv)
    torch.backends.cudnn.benchmark = not a.no_cudnn_benchmark
    tf_peak, hbm, src = peaks()
    net = models.create_model(a.model, kernels="b200")
    layers = OrderedDict()
    for m in net.modules():
        if isinstance(m, Conv):
            key = (m.cin, m.in_hw, m.cout, m.k, m.stride, m.pad)
            layers.setdefault(key, [m, 0])[1] += 1
    rows = []
    tot = {"mine": 0.0, "lib": 0.0}
    print(f"# {a.model} batch {a.batch}; peaks ({src}): {tf_peak:.0f} TFLOP/s bf16, {hbm:.0f} GB/s HBM")
    print(f"{'layer':44s} {'x':>3s} | {'fprop us':>9s} {'lib':>8s} | {'dgrad us':>9s} {'lib':>8s} | {'wgrad us':>9s} {'lib':>8s} |"
          f" {'TF/s f/d/w':>20s} | {'roof us':>8s}")
    for key, (m, count) in layers.items():
        cin, (h, w), cout, k, stride, pad = key
        n = a.batch
        x = torch.randn(n, h, w, cin, device="cuda").bfloat16().permute(0, 3, 1, 2)
        wt = torch.randn(cout, k[0], k[1], cin, device="cuda").bfloat16().permute(0, 3, 1, 2)
        p, q = m.out_hw
        dy = torch.randn(n, p, q, cout, device="cuda").bfloat16().permute(0, 3, 1, 2)
        dw = torch.empty(cout, k[0], k[1], cin, device="cuda").permute(0, 3, 1, 2)
        flops = 2.0 * n * p * q * cout * cin * k[0] * k[1]
        bytes_f = 2.0 * (x.numel() + dy.numel() + wt.numel())
        roof = max(flops / (tf_peak * 1e12), bytes_f / (hbm * 1e9)) * 1e6
        r = {"layer": f"{cin}x{h}x{w}->{cout} k{k[0]}x{k[1]} s{stride[0]}", "count": count, "flops": flops, "roof_us": roof}
        os.environ["B200_CONV"] = "auto"
        if conv_umma.supports_fprop(x, wt, m):
            r["fprop"] = timeit(lambda: conv_umma.fprop(x, wt, m, want_stats=True))
        if conv_umma.supports_dgrad(dy, wt, m, x.shape):
            r["dgrad"] = timeit(lambda: conv_umma.dgrad(dy, wt, m, x.shape))
        if conv_umma.supports_wgrad(x, dy, m):
            r["wgrad"] = timeit(lambda: conv_umma.wgrad(x, dy, m, out=dw))
        if not a.no_lib:
            os.environ["B200_CONV"] = "cudnn"
            r["fprop_lib"] = timeit(lambda: conv_impl.fprop(x, wt, m))
            if cin > 3:
                r["dgrad_lib"] = timeit(lambda: conv_impl.dgrad(dy, wt, m, x.shape))
            r["wgrad_lib"] = timeit(lambda: conv_impl.wgrad(x, dy, m, out=dw))
            os.environ["B200_CONV"] = "auto"
        rows.append(r)

        def f(v):
            return f"{v:9.1f}" if v is not None else "       --"
        tfs = "/".join(f"{flops / (r[kk] * 1e-6) / 1e12:5.0f}" if kk in r else "   --" for kk in ("fprop", "dgrad", "wgrad"))
        print(f"{r['layer']:44s} {count:3d} | {f(r.get('fprop'))} {f(r.get('fprop_lib'))[1:]} | {f(r.get('dgrad'))} "
              f"{f(r.get('dgrad_lib'))[1:]} | {f(r.get('wgrad'))} {f(r.get('wgrad_lib'))[1:]} | {tfs:>20s} | {roof:8.1f}")
        for kk in ("fprop", "dgrad", "wgrad"):
            lib = r.get(kk + "_lib")
            mine = r.get(kk, lib)
            if mine is not None:
                tot["mine"] += mine * count
            if lib is not None:
                tot["lib"] += lib * count
        del x, wt, dy, dw
    print(f"# conv total per step: tcgen05 path {tot['mine'] / 1e3:.2f} ms, library {tot['lib'] / 1e3:.2f} ms")
    if a.json:
        with open(a.json, "w") as fjs:
            json.dump({"model": a.model, "batch": a.batch, "rows": rows, "total_ms": {k: v / 1e3 for k, v in tot.items()}}, fjs, indent=1)
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
