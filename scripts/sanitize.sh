#!/bin/bash
# compute-sanitizer targets for the kernel library (the reference has no sanitizers; its only related
# knob is HOROVOD_MPI_THREADS_DISABLE=1, run-tf-sing-ucx-openmpi.sh:106). One tool per invocation (see
# /opt/skills/guides/B200_PROFILING.md): usage  scripts/sanitize.sh memcheck|racecheck|synccheck|initcheck [pytest -k expr]
set -e
TOOL=${1:-memcheck}
EXPR=${2:-"bn or pool or softmax or normalize or sgd"}
cd "$(dirname "$0")/.."
exec compute-sanitizer --tool "$TOOL" --error-exitcode 1 --launch-timeout 0 \
  python -m pytest tests/test_gpu_kernels.py -x -q -k "$EXPR"
