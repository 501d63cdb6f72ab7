#!/bin/bash
# ncu --set full capture of the two line kernels of one fused 8K run (after the same command exits 0 plainly).
mkdir -p gpurun_out
python tools/prof_cmd.py ${WL:-8k} 3 > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:line_kernel -s 2 -c 2 -f -o gpurun_out/${OUT:-prof_8k} \
    python tools/prof_cmd.py ${WL:-8k} 3 > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
