#!/bin/bash
mkdir -p gpurun_out/q
for shape in "4096 4096 50" "2048 2048 10" "4096 4096 10"; do
  set -- $shape
  for mode in auto off; do
    if [ "$mode" = "off" ]; then export GBLUR_ROW_OLS=0; else unset GBLUR_ROW_OLS; fi
    STEP_CH=1 STEP_ROWS=$1 STEP_COLS=$2 STEP_SIGMA=$3 timeout 200 python tools/step_time.py 2>/dev/null | tee -a gpurun_out/q/gates_c1.jsonl
  done
done
