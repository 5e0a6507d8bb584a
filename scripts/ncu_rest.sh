#!/bin/bash
set -e
OUT=gpurun_out/ncu; mkdir -p $OUT
python scripts/ncu_step.py > $OUT/plain.log 2>&1
ncu --set full --clock-control none --import-source on -k "regex:col_reduce_kernel" -s 10 -c 1 -f -o $OUT/col_reduce_bwd python scripts/ncu_step.py > $OUT/col_reduce_bwd.log 2>&1 || echo fail colreduce
B="--kernel-name-base demangled"
ncu --set full --clock-control none --import-source on $B -k "regex:conv_gemm_kernel<.int.1, .bool.1>" -s 1 -c 1 -f -o $OUT/conv_gemm_halo python scripts/ncu_step.py > $OUT/conv_gemm_halo.log 2>&1 || echo fail halo
ncu --set full --clock-control none --import-source on $B -k "regex:conv_gemm_kernel<.int.1, .bool.0>" -s 6 -c 1 -f -o $OUT/conv_gemm_single python scripts/ncu_step.py > $OUT/conv_gemm_single.log 2>&1 || echo fail single
ncu --set full --clock-control none --import-source on $B -k "regex:conv_gemm_kernel<.int.2, .bool.0>" -s 6 -c 1 -f -o $OUT/conv_gemm_pair python scripts/ncu_step.py > $OUT/conv_gemm_pair.log 2>&1 || echo fail pair
ls -la $OUT/*.ncu-rep | awk '{print $5, $9}' | grep -E "col_reduce|conv_gemm"
tail -2 $OUT/conv_gemm_pair.log
