#!/bin/bash
# Round-2 iteration call: GPU parity tests, the OLS check binary (timings of both passes, parity against the
# whole-line kernels) and one bench line.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
tools/ols_gpu_check 4320 7680 20 auto ${OLS_COL:-0} > gpurun_out/ols_check.jsonl 2>&1; cat gpurun_out/ols_check.jsonl
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json | cut -c1-1500
