#!/bin/bash
# are the overlap-save passes worth it below the automatic rule's gates (rows >= 4096 pixels wide, columns >= 2048 long)?
mkdir -p gpurun_out/q
for shape in "2160 3840 10" "2160 3840 5" "2160 3840 20" "1440 2560 10" "1080 1920 5"; do
  set -- $shape
  for mode in auto forced; do
    if [ "$mode" = "forced" ]; then export GBLUR_ROW_OLS=1 GBLUR_COL_OLS=1; else unset GBLUR_ROW_OLS GBLUR_COL_OLS; fi
    STEP_ROWS=$1 STEP_COLS=$2 STEP_SIGMA=$3 timeout 200 python tools/step_time.py 2>/dev/null | tee -a gpurun_out/q/gates.jsonl
  done
done
