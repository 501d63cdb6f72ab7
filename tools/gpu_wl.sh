#!/bin/bash
# 1-GPU: new tests + the video and band workloads
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "band_sharded_binding or cluster" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
for wl in video_1080p_x512_sigma10 band_16k_gray_sigma50 band_32k_gray_sigma50; do
  timeout 600 python bench.py --workload $wl --steps 5 --warmup 3 > gpurun_out/bench_$wl.json 2> gpurun_out/bench_$wl.err; echo "$wl rc=$?"; cat gpurun_out/bench_$wl.json; tail -5 gpurun_out/bench_$wl.err
done
