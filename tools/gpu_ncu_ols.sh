#!/bin/bash
# ncu of the overlap-save kernels of the 8K step: launch list, then --set full of the row blocks, the column blocks
# and the small pre-kernels. Uses the Python-free check binary, so each ncu command is a few seconds.
# Summaries: tools/ncu_summary.py on the .ncu-rep files brought back in gpurun_out/.
mkdir -p gpurun_out
tools/ols_gpu_check 4320 7680 20 auto 0 > gpurun_out/plain_ols.log 2>&1 || { cat gpurun_out/plain_ols.log; exit 1; }
cat gpurun_out/plain_ols.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_8k_ols.csv \
    tools/ols_gpu_check 4320 7680 20 auto 0 > gpurun_out/ncu_launches_ols.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:ols_row_rgba -s 1 -c 1 -f -o gpurun_out/prof_ols_row \
    tools/ols_gpu_check 4320 7680 20 auto 0 > gpurun_out/ncu_ols_row.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:ols_col_kernel -s 1 -c 1 -f -o gpurun_out/prof_ols_col \
    tools/ols_gpu_check 4320 7680 20 auto 0 > gpurun_out/ncu_ols_col.log 2>&1
ncu --set full --clock-control none -k regex:'altsum_rows_kernel|colsum_kernel|u_rowpass_kernel' -s 6 -c 3 -f -o gpurun_out/prof_ols_pre \
    tools/ols_gpu_check 4320 7680 20 auto 0 > gpurun_out/ncu_ols_pre.log 2>&1
tail -2 gpurun_out/ncu_ols_*.log; ls -la gpurun_out/
