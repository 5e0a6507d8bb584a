#!/bin/bash
# ./run-tf-b200.sh 1 4 256 nvl
# usage: ./run-tf-b200.sh <NUM_NODES> <WORKERS_PER_SOCKET> <batch_size> <fabric(nvl|ib|nccl|sock)> [extra cnn_benchmarks flags]
# If WORKERS_PER_SOCKET=0, then WORKERS_PER_NODE=1
#
# Same four positionals as the reference's benchmark-scripts/run-tf-sing-ucx-openmpi.sh (:27-30) and
# run-tf-sing-libfabric-intelmpi.sh. A "worker" is one process driving one B200; a "socket" is a NUMA node.
# Fabrics: nvl = hand-written NVLink/NVSwitch kernels (ib is accepted as an alias), nccl = stock NCCL
# baseline, sock = gloo over TCP on CPU.
#
# Examples, assuming a 2-socket server with 4 GPUs per socket:
#   8 workers over NVLink:  ./run-tf-b200.sh 1 4 256 nvl  2>&1 | tee ~/logs/tfmn-1n-256b-synth-nvl-r1.log
#   2 CPU workers, sockets: ./run-tf-b200.sh 1 1 32 sock  2>&1 | tee ~/logs/tfmn-1n-32b-synth-sock-r1.log
#
# The constants the reference hard-codes (:32-35) can be overridden from the environment:
#   MODEL=vgg16 NUM_BATCHES=100 NUM_WARMUP_BATCHES=50 INTER_T=2 ./run-tf-b200.sh 1 4 64 nvl
set -e
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
export PYTHONPATH="${HERE}/..:${PYTHONPATH}"
# HOROVOD_FUSION_THRESHOLD (the reference exports 134217728, :105) is honoured when set; unset, the launcher picks
# 128 MiB for nccl/sock and 16 MiB for nvl (buckets that overlap with backward).
exec python -m azure_hc_intel_tf_b200.parallel.launcher "$@"
