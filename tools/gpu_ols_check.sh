#!/bin/bash
# one-shot GPU check of the overlap-save passes (parity vs the whole-line kernels + timings), no Python
# args of tools/ols_gpu_check: rows cols sigma row-variant|auto col-ols(0|1)
mkdir -p gpurun_out
run() { timeout 20 tools/ols_gpu_check "$@" 2>&1 | tee -a gpurun_out/ols_check.jsonl; }
run 4320 7680 20 auto 1
run 1500 4100 7 1 1
run 160 7680 20 1 1
