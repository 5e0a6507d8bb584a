#!/bin/bash
# One `ncu --set full` capture per kernel of the ResNet-50 step (second launch of each, batch 256), reports under
# gpurun_out/ncu/. Run on ONE GPU through gpurun, after `python scripts/ncu_step.py` has exited 0 there.
set -e
OUT=gpurun_out/ncu; mkdir -p $OUT
python scripts/ncu_step.py > $OUT/plain.log 2>&1
for k in bn_apply_kernel bn_bwd_apply_kernel "col_reduce_kernel.*BwdReduce" bn_finalize_partials_kernel add_masked_kernel add_kernel \
         maxpool_fwd maxpool_bwd xent_rows_kernel sgd_flat_kernel im2col_kernel wgrad_fold_kernel spatial_mean_fwd conv_wgrad_kernel normalize_u8; do
  name=$(echo "$k" | tr -cd 'A-Za-z0-9_')
  ncu --set full --clock-control none --import-source on -k "regex:$k" -s 1 -c 1 -f -o $OUT/$name python scripts/ncu_step.py > $OUT/$name.log 2>&1 || echo "ncu failed for $k"
done
# the three conv_gemm variants: halo <1,1>, single-CTA <1,0>, CTA pair <2,0>
ncu --set full --clock-control none --import-source on -k "regex:conv_gemm_kernel<1, true>" -s 1 -c 1 -f -o $OUT/conv_gemm_halo python scripts/ncu_step.py > $OUT/conv_gemm_halo.log 2>&1 || echo fail halo
ncu --set full --clock-control none --import-source on -k "regex:conv_gemm_kernel<1, false>" -s 6 -c 1 -f -o $OUT/conv_gemm_single python scripts/ncu_step.py > $OUT/conv_gemm_single.log 2>&1 || echo fail single
ncu --set full --clock-control none --import-source on -k "regex:conv_gemm_kernel<2, false>" -s 6 -c 1 -f -o $OUT/conv_gemm_pair python scripts/ncu_step.py > $OUT/conv_gemm_pair.log 2>&1 || echo fail pair
ls -la $OUT/*.ncu-rep | awk '{print $5, $9}'
