#!/bin/bash
mkdir -p gpurun_out/q
timeout 600 python -m pytest tests/test_zz_no_alloc.py tests/test_zz_gpu_halo.py -m gpu -x -q > gpurun_out/q/pytest_colsum.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/q/pytest_colsum.log
for r in 64 32 16 128; do
  GBLUR_COLSUM_ROWS=$r timeout 300 python tools/step_time.py 2>/dev/null | tee -a gpurun_out/q/colsum_rows_groups4.jsonl
done
