#!/bin/bash
# Round-2 first GPU call: fp32 issue microbenchmark (scalar vs packed f32x2), measure the unmeasured OLS variants,
# then ncu of the overlap-save kernels as they were at the end of round 1.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt
tools/ubench/ubench_fp32 > gpurun_out/ubench_fp32.jsonl 2>&1; echo "ubench rc=$?"
for v in 1 5; do tools/ols_gpu_check 4320 7680 20 $v 0 >> gpurun_out/ols_variants.jsonl 2>&1; done
GBLUR_OLS_COL_VARIANT=1 tools/ols_gpu_check 4320 7680 20 auto 1 >> gpurun_out/ols_variants.jsonl 2>&1
bash tools/gpu_ncu_ols.sh > gpurun_out/ncu_ols_script.log 2>&1
tail -40 gpurun_out/ubench_fp32.jsonl; cat gpurun_out/ols_variants.jsonl
