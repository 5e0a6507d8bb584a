#!/bin/bash
# Trimmed whole-box check (N GPUs): comm correctness tests, the headline bench, the per-algorithm allreduce sweep.
N=${1:-8}; O=gpurun_out/n$N; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 400 python -m pytest tests/test_gpu_comm.py -m gpu -q -k "large_buckets or own_kernels or p2p_allreduce" > $O/pytest_comm.log 2>&1; echo "pytest rc=$?" >> $O/pytest_comm.log; tail -3 $O/pytest_comm.log
timeout 400 $TR --master-port 29701 bench.py --gpus $N --steps 20 --warmup 5 --no-e2e > $O/bench_resnet50.json 2> $O/bench_resnet50.err; echo "bench rc=$?"; cat $O/bench_resnet50.json; tail -2 $O/bench_resnet50.err
timeout 300 $TR --master-port 29705 -m azure_hc_intel_tf_b200.bench.allreduce_sweep --op allreduce --algos auto,oneshot,twoshot,nvls --dtypes float32 --trials 3 --max-bytes 268435456 --json $O/allreduce_${N}gpu.json 2>&1 | grep -v "^\*\|OMP_NUM\|^$" | tee $O/allreduce_${N}gpu.txt
