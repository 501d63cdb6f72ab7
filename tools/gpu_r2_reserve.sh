#!/bin/bash
mkdir -p gpurun_out/ev
timeout 900 python -m pytest tests/test_zz_gpu_halo.py tests/test_zz_gpu_ols_rgb.py -m gpu -x -q > gpurun_out/ev/pytest_halo_rgb.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/ev/pytest_halo_rgb.log
for r in 0 1 2 4 8; do
  GBLUR_ROW_RESERVE_SMS=$r timeout 300 python tools/step_time.py 2>/dev/null | tee -a gpurun_out/ev/row_reserve.jsonl
done
