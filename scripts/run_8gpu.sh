#!/bin/bash
# Everything that needs the whole box, in one gpurun call (N = number of GPUs, default 8).
N=${1:-8}; O=gpurun_out/n$N; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_gpu_comm.py -m gpu -q > $O/pytest_comm.log 2>&1; echo "pytest rc=$?" >> $O/pytest_comm.log; tail -3 $O/pytest_comm.log
timeout 600 $TR --master-port 29701 bench.py --gpus $N --steps 30 --warmup 5 > $O/bench_resnet50.json 2> $O/bench_resnet50.err; echo "bench rc=$?"; cat $O/bench_resnet50.json
timeout 300 $TR --master-port 29702 bench.py --gpus $N --steps 20 --warmup 5 --model vgg16 --batch 64 --no-e2e > $O/bench_vgg16.json 2> $O/bench_vgg16.err; cat $O/bench_vgg16.json
timeout 300 $TR --master-port 29703 bench.py --gpus $N --steps 20 --warmup 5 --optimizer lars --no-e2e --no-baseline > $O/bench_resnet50_lars.json 2> $O/bench_lars.err; cat $O/bench_resnet50_lars.json
timeout 300 $TR --master-port 29704 -m azure_hc_intel_tf_b200.cli.cnn_benchmarks --model=resnet50 --batch_size=256 --num_batches=10 --num_warmup_batches=5 --variable_update=horovod --fabric=nvl --kernels=b200 --device=gpu --trace_file=$O/trace_step.json > $O/trace.log 2>&1; tail -4 $O/trace.log
timeout 900 bash scripts/osu_suite.sh $N $O/osu > $O/osu.log 2>&1; grep -c . $O/osu.log
