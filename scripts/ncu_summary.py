"""Summarise an .ncu-rep (read here without a GPU) into a small markdown file for profiles/."""
import csv, subprocess, sys
rep, out = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__shared_mem_per_block_dynamic", "sm__cycles_elapsed.avg.per_second",
        "sm__inst_executed_pipe_tensor_op_hmma.sum", "smsp__inst_executed.sum"]
with open(out, "w") as f:
    f.write(f"# ncu summary of `{rep.split('/')[-1]}` (--set full --clock-control none)\n\n")
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")]
        f.write(f"## {name[:120]}\n\n| metric | value | unit |\n|---|---|---|\n")
        for w in want:
            if w in hdr:
                i = hdr.index(w)
                f.write(f"| {w} | {r[i]} | {units[i]} |\n")
        f.write("\n")
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    srows = list(csv.reader(src.splitlines()))
    if len(srows) > 2:
        h = srows[1]
        if "Warp Stall Sampling (All Samples)" in h:
            si = h.index("Warp Stall Sampling (All Samples)")
            data = [(int(r[si]), r[1].strip()) for r in srows[2:] if len(r) > si and r[si].isdigit()]
            tot = sum(d[0] for d in data) or 1
            f.write("## hottest SASS lines by warp-stall samples (last kernel in the report)\n\n```\n")
            for s, ins in sorted(data, reverse=True)[:12]:
                f.write(f"{s:6d} {100 * s / tot:5.1f}%  {ins[:110]}\n")
            f.write("```\nSpin loops on mbarrier try_wait dominate the samples by construction: producer, MMA-issuer and "
                    "epilogue warps wait on each other; tensor-pipe and memory percentages above are the signal.\n")
print("wrote", out)
