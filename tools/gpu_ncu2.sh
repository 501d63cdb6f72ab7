#!/bin/bash
# ncu --set full of one cluster kernel (config-4 shaped rows / columns): WL=c4row|c4col
mkdir -p gpurun_out
wl=${WL:-c4row}
python tools/prof_cmd.py $wl 2 > gpurun_out/plain_$wl.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:line_kernel_cluster -s 1 -c 1 -f -o gpurun_out/prof_$wl \
    python tools/prof_cmd.py $wl 2 > gpurun_out/ncu_$wl.log 2>&1
tail -2 gpurun_out/ncu_$wl.log; ls -la gpurun_out/
