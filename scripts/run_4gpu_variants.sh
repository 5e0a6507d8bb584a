O=gpurun_out/n4; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
timeout 400 python -m pytest tests/test_gpu_comm.py -m gpu -q -k "large_buckets or p2p_allreduce or training" > $O/pytest_comm.log 2>&1; echo "pytest rc=$?" >> $O/pytest_comm.log; tail -3 $O/pytest_comm.log
run() { timeout 200 $TR --master-port $1 bench.py --gpus 4 --steps 20 --warmup 5 --no-e2e --no-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$2', d['value'], d['ms_per_step'], d['config']['fusion_buckets'])"; }
run 29711 default
B200_COMM_CTAS=32 run 29712 ctas32
HOROVOD_FUSION_THRESHOLD=33554432 run 29713 thr32M
HOROVOD_FUSION_THRESHOLD=8388608 run 29714 thr8M
