#!/bin/bash
# one-shot GPU check of the overlap-save row pass (parity vs the whole-line kernels + timings), no Python
mkdir -p gpurun_out
run() { timeout 40 tools/ols_gpu_check "$@" 2>&1 | tee -a gpurun_out/ols_check.jsonl; }
run 4320 7680 20 auto
run 160 7680 20 auto
run 2000 5000 50 auto
run 600 9000 2 auto
run 300 4100 7 1
run 133 1203 20 1
run 4320 7680 20 4
