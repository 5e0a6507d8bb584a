#!/bin/bash
# Runs the whole OSU-role suite at N GPUs and writes one JSON per (op, N) under profiles/osu/ (via gpurun_out/ when run
# through gpurun). usage: scripts/osu_suite.sh <N> [out_dir] [max_bytes]
set -e
N=${1:-2}; OUT=${2:-gpurun_out/osu}; MAX=${3:-1073741824}
mkdir -p "$OUT"
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) \
        -m azure_hc_intel_tf_b200.bench.allreduce_sweep "$@"; }
run --op allreduce --algos auto,oneshot,twoshot,nvls --max-bytes "$MAX" --json "$OUT/allreduce_${N}gpu.json" 2>&1 | grep -v "^\*\|OMP_NUM" | tee "$OUT/allreduce_${N}gpu.txt"
run --op allgather --max-bytes "$MAX" --json "$OUT/allgather_${N}gpu.json" 2>&1 | grep -v "^\*\|OMP_NUM" | tee "$OUT/allgather_${N}gpu.txt"
run --op broadcast --max-bytes "$MAX" --json "$OUT/broadcast_${N}gpu.json" 2>&1 | grep -v "^\*\|OMP_NUM" | tee "$OUT/broadcast_${N}gpu.txt"
run --op barrier --json "$OUT/barrier_${N}gpu.json" 2>&1 | grep -v "^\*\|OMP_NUM" | tee "$OUT/barrier_${N}gpu.txt"
run --op p2p --max-bytes 268435456 --json "$OUT/p2p_${N}gpu.json" 2>&1 | grep -v "^\*\|OMP_NUM" | tee "$OUT/p2p_${N}gpu.txt"
