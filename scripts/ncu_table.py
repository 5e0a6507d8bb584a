"""Summarise every .ncu-rep under gpurun_out/ncu/ into profiles/ncu_step_kernels.md (one row per captured kernel) and one
profiles/ncu/<name>.md each (scripts/ncu_summary.py format). Runs here, without a GPU."""
import csv, glob, json, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "gpurun_out", "ncu")
try:
    PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
except Exception:
    PEAK = {"hbm_gbs": 6650.0}
HBM = PEAK["hbm_gbs"]
rows_out = []
os.makedirs(os.path.join(ROOT, "profiles", "ncu"), exist_ok=True)
for rep in sorted(glob.glob(os.path.join(SRC, "*.ncu-rep"))):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    if len(rows) < 3:
        continue
    h, u, r = rows[0], rows[1], rows[-1]
    def g(key, default=0.0):
        if key in h:
            try:
                return float(r[h.index(key)].replace(",", ""))
            except ValueError:
                return default
        return default
    def unit(key):
        return u[h.index(key)] if key in h else ""
    def to_bytes(key):
        v, un = g(key), unit(key).lower()
        return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(un, 1)
    dur = g("gpu__time_duration.sum")
    dur_us = dur * {"ns": 1e-3, "us": 1, "ms": 1e3}.get(unit("gpu__time_duration.sum"), 1)
    rd, wr = to_bytes("dram__bytes_read.sum"), to_bytes("dram__bytes_write.sum")
    gbs = (rd + wr) / (dur_us * 1e-6) / 1e9 if dur_us else 0
    name = re.sub(r"void |\(anonymous namespace\)::|unnamed>::|b200::|<unnamed>::", "", r[h.index("Kernel Name")])
    name = re.sub(r"\(.*", "", name)[:70]
    rows_out.append((os.path.basename(rep)[:-8], name, r[h.index("Grid Size")] if "Grid Size" in h else "", r[h.index("Block Size")] if "Block Size" in h else "",
                     dur_us, rd / 1e6, wr / 1e6, gbs, 100 * gbs / HBM,
                     g("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"),
                     g("lts__throughput.avg.pct_of_peak_sustained_elapsed"),
                     g("sm__warps_active.avg.pct_of_peak_sustained_active"), int(g("launch__registers_per_thread"))))
    subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "ncu_summary.py"), rep,
                    os.path.join(ROOT, "profiles", "ncu", os.path.basename(rep)[:-8] + ".md")], capture_output=True)
with open(os.path.join(ROOT, "profiles", "ncu_step_kernels.md"), "w") as f:
    f.write("# ncu captures of the ResNet-50 training step's kernels (B200, batch 256, bf16; `ncu --set full --clock-control none`)\n\n"
            "One launch per kernel (scripts/ncu_all.sh, scripts/ncu_rest.sh; per-kernel detail and hottest SASS lines in profiles/ncu/). "
            f"DRAM GB/s = (dram bytes read + written) / duration; `% of copy` is against the measured copy bandwidth {HBM:.0f} GB/s "
            "(MEASURED_PEAKS.json). Durations under ncu are cold-cache single launches: compare shares, not absolutes.\n\n"
            "| capture | kernel | grid | block | µs | DRAM rd MB | DRAM wr MB | DRAM GB/s | % of copy (measured) | tensor pipe % | L2 % | warps active % | regs |\n"
            "|---|---|---|---|---:|---:|---:|---:|---:|---:|---:|---:|---:|\n")
    for t in rows_out:
        f.write("| %s | `%s` | %s | %s | %.1f | %.1f | %.1f | %.0f | %.0f | %.1f | %.1f | %.1f | %d |\n" % t)
print("wrote", len(rows_out), "rows")
