#!/bin/bash
# GPU check of the overlap-save passes without Python: parity against the whole-line kernels and CUDA-event timings.
# args of tools/ols_gpu_check (build: tools/build_ols_check.sh): rows cols sigma row-variant|auto col-ols(0|1)
# GBLUR_OLS_COL_VARIANT: 0 = four warps per quarter panel, 1 = eight warps per half panel
mkdir -p gpurun_out
run() { timeout 20 tools/ols_gpu_check "$@" 2>&1 | tee -a gpurun_out/ols_check.jsonl; }
run 4320 7680 20 auto 0
run 4320 7680 20 auto 1
GBLUR_OLS_COL_VARIANT=1 run 4320 7680 20 auto 1
run 1500 4100 7 1 1
GBLUR_OLS_COL_VARIANT=1 run 1500 4100 7 1 1
run 160 7680 20 1 1
# row variants: 5 = one CTA per row, alternating sums inside the block kernel (no pre-kernel)
run 4320 7680 20 5 0
run 1500 4100 7 5 0
