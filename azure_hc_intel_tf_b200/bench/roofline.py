"""Achieved fraction of roofline for every fused path, from measured files (no GPU needed).

BASELINE.json's north star: "each fused path is reported as achieved fraction of its roofline (the slower of its
compute at peak and its bytes over NVLink at link bandwidth)". Denominators (profiling guide + MEASURED_PEAKS.json):
measured copy bandwidth (HBM), measured sustained cuBLAS bf16 rate, and for NVLink the measured peer-copy figure
(profiles/osu/p2p_*: this tool takes the best push bandwidth found there, else the guide's 770 GB/s per direction).

    python -m azure_hc_intel_tf_b200.bench.roofline [--conv profiles/kernel_bench_*.json] [--osu profiles/osu] [--md out.md]

Paths reported
* convolution fprop / dgrad / wgrad per layer: time vs max(FLOPs / peak, bytes / HBM)  (from bench/kernel_bench.py);
* allreduce per size and rank count: per GPU and direction the link carries (N-1)/N*S (two-shot: reduce-scatter in,
  all-gather out) or (1 + 1/N)*S... no: S/N in + S out for NVLS multimem (the switch reduces and replicates) - see
  link_bytes(); one-shot carries (N-1)*S in;
* fused allreduce + momentum-SGD per bucket: max(link bytes / NVLink, HBM bytes of g/w/m/shadow / HBM). The fused
  kernel's time is taken as the allreduce time of that size (the update half is the HBM term, hidden under it when the
  link term dominates); measured end to end by the scaling runs in BASELINE.md.

Reference role: none (the reference publishes no roofline); this is the product definition's reporting requirement.
"""
from __future__ import annotations

import argparse
import glob
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return {"tflops": p.get("bf16_tflops_sustained", p["bf16_tflops"]), "hbm": p["hbm_gbs"], "src": "measured"}
    except Exception:
        return {"tflops": 1590.0, "hbm": 6650.0, "src": "fallback"}


def nvlink_gbs(osu_dir: str) -> tuple:
    best = 0.0
    for fn in glob.glob(os.path.join(osu_dir, "p2p_*gpu.json")):
        for r in json.load(open(fn)).get("rows", []):
            best = max(best, r.get("push_gbs", 0.0), r.get("pull_gbs", 0.0))
    return (best, "measured peer copy") if best > 0 else (770.0, "profiling guide")


def link_bytes(algo: str, size: float, n: int) -> float:
    """Bytes that cross one GPU's NVLink in its busier direction for an allreduce of `size` bytes over n ranks."""
    if algo == "oneshot":
        return (n - 1) * size                       # every rank pulls every peer's copy
    if algo == "nvls":
        # out: this GPU's copy of every shard goes up to the switch once (S) + its reduced shard for the broadcast (S/N);
        # in: its reduced shard (S/N) + everyone's shards replicated (S)
        return size * (1.0 + 1.0 / n)
    return size * (n - 1) / n                       # two-shot: (N-1)/N in for the reduce, the same out for the gather


def conv_rows(path: str, pk: dict):
    d = json.load(open(path))
    out = []
    for r in d["rows"]:
        for op in ("fprop", "dgrad", "wgrad"):
            if op in r:
                out.append((r["layer"], op, r[op], r["roof_us"], r["roof_us"] / r[op], r.get(op + "_lib")))
    return d.get("model", "?"), out


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--conv", default=None, help="kernel_bench JSON (default: newest profiles/kernel_bench_*.json)")
    ap.add_argument("--osu", default=os.path.join(ROOT, "profiles", "osu"))
    ap.add_argument("--md", default=None)
    a = ap.parse_args(argv)
    pk = peaks()
    nvl, nvl_src = nvlink_gbs(a.osu)
    lines = [f"# Roofline fractions (denominators: {pk['tflops']:.0f} TFLOP/s bf16 sustained and {pk['hbm']:.0f} GB/s copy "
             f"({pk['src']}); NVLink {nvl:.0f} GB/s per direction ({nvl_src}))", ""]
    conv = a.conv or (sorted(glob.glob(os.path.join(ROOT, "profiles", "kernel_bench_resnet50*.json")), key=os.path.getmtime) or [None])[-1]
    if conv:
        model, rows = conv_rows(conv, pk)
        lines += [f"## Convolution kernels, {model} ({os.path.basename(conv)})", "",
                  "| layer | op | µs | roofline µs | achieved fraction | library µs |", "|---|---|---:|---:|---:|---:|"]
        tot = {}
        for layer, op, t, roof, frac, lib in rows:
            lines.append(f"| {layer} | {op} | {t:.1f} | {roof:.1f} | {frac:.2f} | {lib:.1f} |" if lib else
                         f"| {layer} | {op} | {t:.1f} | {roof:.1f} | {frac:.2f} | -- |")
            s = tot.setdefault(op, [0.0, 0.0])
            s[0] += t
            s[1] += roof
        lines += [""] + [f"* {op}: sum of distinct layers {v[0]:.0f} µs vs roofline {v[1]:.0f} µs = {v[1] / v[0]:.2f}" for op, v in tot.items()] + [""]
    for fn in sorted(glob.glob(os.path.join(a.osu, "allreduce_*gpu.json"))):
        d = json.load(open(fn))
        n = d["ranks"]
        lines += [f"## Allreduce and fused allreduce+SGD, {n} GPUs ({os.path.basename(fn)})", "",
                  "| bytes | algorithm | µs | link bytes/dir MB | link-bound µs | allreduce fraction | fused: HBM-bound µs | fused roofline µs | fused fraction |",
                  "|---:|---|---:|---:|---:|---:|---:|---:|---:|"]
        for r in d["rows"]:
            if r["dtype"] != "float32" or r["bytes"] < (1 << 20):
                continue
            cands = {k[:-3]: v for k, v in r.items() if k.endswith("_us") and k[:-3] in ("oneshot", "twoshot", "nvls")}
            if not cands:
                continue
            algo = min(cands, key=cands.get)
            t = cands[algo]
            lb = link_bytes(algo, r["bytes"], n)
            t_link = lb / (nvl * 1e3)
            # fused update per parameter: read g (4) w (4) m (4), write w (4) m (4) bf16 shadow (2) = 22 B per 4 B of message,
            # plus the reduce's own local traffic (read own copy, write reduced result): 8 B
            t_hbm = r["bytes"] / 4 * 30 / (pk["hbm"] * 1e3)
            roof = max(t_link, t_hbm)
            lines.append(f"| {r['bytes']} | {algo} | {t:.1f} | {lb / 1e6:.1f} | {t_link:.1f} | {t_link / t:.2f} | {t_hbm:.1f} | {roof:.1f} | "
                         f"{roof / max(t, roof):.2f} |")
        lines.append("")
    text = "\n".join(lines)
    print(text)
    if a.md:
        with open(a.md, "w") as f:
            f.write(text + "\n")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
