"""Time the BN backward column reduction (b200_bn_bwd_reduce) on the ResNet-50 activation shapes at batch 256, CUDA events,
median of 20 graph replays of 10 calls each, against the copy roofline (bytes of dz + y [+ mask] / measured copy bandwidth).
usage: python scripts/bn_bench.py"""
import json
import os
import sys
import torch
sys.path.insert(0, ".")
from azure_hc_intel_tf_b200.ops.b200 import _check, _ptr, _stream, lib, scratch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
try:
    HBM = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    HBM = 6650.0


def med(fn, n=20, reps=10):
    """Median time of one call: `reps` calls are captured in a CUDA graph and the replay is timed, so that the host's
    launch latency (tens of microseconds through ctypes) is not in the figure."""
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3):
            fn()
    torch.cuda.current_stream().wait_stream(side)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    g.replay()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
    for a, c in ev:
        a.record(); g.replay(); c.record()
    torch.cuda.synchronize()
    t = sorted(a.elapsed_time(c) for a, c in ev)
    return t[n // 2] * 1e3 / reps


dev = torch.device("cuda")
tot = {"mask": [0.0, 0.0], "recompute": [0.0, 0.0]}
print("| rows x c | mode | us | GB/s | % of copy |\n|---|---|---:|---:|---:|")
for hw, c, count in ((56, 64, 6), (56, 256, 4), (28, 128, 8), (28, 512, 5), (14, 256, 12), (14, 1024, 7), (7, 512, 6), (7, 2048, 4)):
    rows = 256 * hw * hw
    dz = torch.randn(rows, c, device=dev).to(torch.bfloat16)
    y = torch.randn(rows, c, device=dev).to(torch.bfloat16)
    mask = torch.randint(0, 256, (rows * c // 8,), dtype=torch.uint8, device=dev)
    st = torch.rand(4, c, device=dev) + 0.5
    dg, db = torch.empty(c, device=dev), torch.empty(c, device=dev)
    for mode, mk in (("mask", mask), ("recompute", None)):
        fn = lambda: _check(lib().b200_bn_bwd_reduce(_ptr(dz), _ptr(y), _ptr(None), _ptr(mk), rows, c, _ptr(st[0]), _ptr(st[1]),
                                                      _ptr(st[2]), _ptr(st[3]), 1, _ptr(dg), _ptr(db), _ptr(scratch(dev)), _stream()),
                            "bn_bwd_reduce")
        t = med(fn)
        byts = rows * c * 4 + (rows * c // 8 if mk is not None else 0)
        gbs = byts / t / 1e3
        tot[mode][0] += t * count
        tot[mode][1] += byts / HBM / 1e3 * count
        print(f"| {rows} x {c} | {mode} | {t:.1f} | {gbs:.0f} | {100 * gbs / HBM:.0f} |")
for mode, (t, r) in tot.items():
    print(f"\nResNet-50 weighted sum, {mode}: {t / 1e3:.2f} ms vs roofline {r / 1e3:.2f} ms ({100 * r / t:.0f} %)")

# the per-layer statistics fold (conv-epilogue partial sums -> mean / rstd / scale / shift): a latency chain, not bandwidth
print("\n| partial rows x c | bn_finalize_partials us |\n|---|---:|")
fin = 0.0
for nrows, c, count in ((148, 64, 7), (148, 256, 16), (148, 128, 8), (148, 512, 11), (148, 1024, 7), (148, 2048, 4), (296, 256, 4)):
    part = torch.randn(nrows, 2, c, device=dev).abs()
    g_, b_ = torch.ones(c, device=dev), torch.zeros(c, device=dev)
    rm, rv = torch.zeros(c, device=dev), torch.ones(c, device=dev)
    out = torch.empty(4, c, device=dev)
    fn = lambda: _check(lib().b200_bn_finalize_partials(_ptr(part), nrows, 200704, c, _ptr(g_), _ptr(b_), _ptr(rm), _ptr(rv), 0.9, 1e-5,
                                                        _ptr(out[0]), _ptr(out[1]), _ptr(out[2]), _ptr(out[3]), _stream()), "fin")
    t = med(fn)
    fin += t * count
    print(f"| {nrows} x {c} | {t:.1f} |")
print(f"\nResNet-50 weighted sum of the statistics folds: {fin / 1e3:.2f} ms")
