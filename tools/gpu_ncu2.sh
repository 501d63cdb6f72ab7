#!/bin/bash
# ncu --set full of the cluster kernels (config-4 shaped rows / columns)
mkdir -p gpurun_out
for wl in c4row c4col; do
python tools/prof_cmd.py $wl 2 > gpurun_out/plain_$wl.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:line_kernel -s 2 -c 2 -f -o gpurun_out/prof_$wl \
    python tools/prof_cmd.py $wl 2 > gpurun_out/ncu_$wl.log 2>&1
tail -2 gpurun_out/ncu_$wl.log
done
