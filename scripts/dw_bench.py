"""Time the depthwise kernels (csrc/kernels/depthwise.cu) on the MobileNet-v2 layer shapes against their HBM roofline
(bytes moved / measured copy bandwidth) and against the library (cuDNN through aten::convolution), CUDA events, L2
flushed between iterations. Writes profiles/depthwise_bench.json + .md.

    python scripts/dw_bench.py [--batch 64] [--iters 20]
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from azure_hc_intel_tf_b200.models import create_model                      # noqa: E402
from azure_hc_intel_tf_b200.models.builder import Conv                      # noqa: E402
from azure_hc_intel_tf_b200.ops import conv_umma                            # noqa: E402


def timed(fn, iters, flush):
    for _ in range(3):
        fn()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for a, b in ev:
        flush.add_(1.0)                                                     # 256 MiB write: evicts the 126 MB L2
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    t = sorted(a.elapsed_time(b) for a, b in ev)
    return t[len(t) // 2] * 1e3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles"))
    a = ap.parse_args()
    try:
        hbm = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        hbm = 6650.0
    torch.backends.cudnn.benchmark = True
    flush = torch.zeros(64 << 20, device="cuda")
    model = create_model("mobilenet", kernels="b200")
    seen, rows = set(), []
    for m in model.modules():
        if not (isinstance(m, Conv) and m.groups > 1):
            continue
        key = (m.cin, m.in_hw, m.stride)
        if key in seen:
            continue
        seen.add(key)
        n, c, (h, w), (p, q) = a.batch, m.cin, m.in_hw, m.out_hw
        x = torch.randn(n, h, w, c, device="cuda").to(torch.bfloat16).permute(0, 3, 1, 2)
        dy = torch.randn(n, p, q, c, device="cuda").to(torch.bfloat16).permute(0, 3, 1, 2)
        wt = torch.randn(c, 1, 3, 3, device="cuda").to(torch.bfloat16)
        (pt, pb), (pl, pr) = m.pad
        xp = torch.nn.functional.pad(x, (pl, pr, pt, pb)).contiguous(memory_format=torch.channels_last)
        in_b, out_b = x.numel() * 2, dy.numel() * 2
        row = {"layer": f"{c}x{h}x{w}/s{m.stride[0]}", "roof_fprop_us": (in_b + out_b) / hbm / 1e3,
               "roof_wgrad_us": (in_b + out_b) / hbm / 1e3}
        row["fprop"] = timed(lambda: conv_umma.dw_fprop(x, wt, m), a.iters, flush)
        row["dgrad"] = timed(lambda: conv_umma.dw_dgrad(dy, wt, m, x.shape), a.iters, flush)
        row["wgrad"] = timed(lambda: conv_umma.dw_wgrad(x, dy, m), a.iters, flush)
        conv = torch.ops.aten.convolution
        bwd = torch.ops.aten.convolution_backward
        row["fprop_lib"] = timed(lambda: conv(xp, wt, None, m.stride, (0, 0), (1, 1), False, (0, 0), c), a.iters, flush)
        row["dgrad_lib"] = timed(lambda: bwd(dy, xp, wt, None, m.stride, (0, 0), (1, 1), False, (0, 0), c, [True, False, False]),
                                 a.iters, flush)
        row["wgrad_lib"] = timed(lambda: bwd(dy, xp, wt, None, m.stride, (0, 0), (1, 1), False, (0, 0), c, [False, True, False]),
                                 a.iters, flush)
        rows.append(row)
        print(row, flush=True)
    out = {"model": "mobilenet", "batch": a.batch, "hbm_gbs": hbm, "l2": "flushed between iterations (256 MiB write)", "rows": rows}
    os.makedirs(a.out, exist_ok=True)
    json.dump(out, open(os.path.join(a.out, "depthwise_bench.json"), "w"), indent=1)
    with open(os.path.join(a.out, "depthwise_bench.md"), "w") as f:
        f.write(f"# Depthwise 3x3 kernels, MobileNet-v2 shapes, batch {a.batch} (µs, median of {a.iters}, L2 flushed; roofline = "
                f"(input + output bytes) / {hbm:.0f} GB/s). About 13 µs of every figure, the library's included, is a floor of this harness: "
                "the host's launch latency behind the flush kernel and the write-back of the flush's dirty L2 lines)\n\n"
                "| layer | roofline | fprop | dgrad | wgrad | cuDNN fprop | cuDNN dgrad | cuDNN wgrad |\n"
                "|---|---:|---:|---:|---:|---:|---:|---:|\n")
        for r in rows:
            f.write(f"| {r['layer']} | {r['roof_fprop_us']:.1f} | {r['fprop']:.1f} | {r['dgrad']:.1f} | {r['wgrad']:.1f} | "
                    f"{r['fprop_lib']:.1f} | {r['dgrad_lib']:.1f} | {r['wgrad_lib']:.1f} |\n")
        tot = {k: sum(r[k] for r in rows) for k in ("roof_fprop_us", "fprop", "dgrad", "wgrad", "fprop_lib", "dgrad_lib", "wgrad_lib")}
        f.write(f"| **sum** | {tot['roof_fprop_us']:.0f} | {tot['fprop']:.0f} | {tot['dgrad']:.0f} | {tot['wgrad']:.0f} | "
                f"{tot['fprop_lib']:.0f} | {tot['dgrad_lib']:.0f} | {tot['wgrad_lib']:.0f} |\n")


if __name__ == "__main__":
    main()
