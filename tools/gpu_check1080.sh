#!/bin/bash
# Parity suite + headline bench + 1080p video bench + row/column phase split (no profiler).
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json
python bench.py --workload video_1080p_x512_sigma10 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_video_1080p_x512_sigma10.json 2> gpurun_out/bench_video_1080p_x512_sigma10.err; echo "video rc=$?"; cat gpurun_out/bench_video_1080p_x512_sigma10.json
python bench.py --workload 1080p_rgb_sigma5 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_1080p_rgb_sigma5.json 2> gpurun_out/bench_1080p_rgb_sigma5.err; echo "1080p rc=$?"; cat gpurun_out/bench_1080p_rgb_sigma5.json
python tools/phase_times.py 1080 1920 3 10 0 16 > gpurun_out/phase_1080p_x16.json 2>&1; cat gpurun_out/phase_1080p_x16.json
python tools/phase_times.py 1080 1920 4 10 1 16 > gpurun_out/phase_1080p_rgba_x16.json 2>&1; cat gpurun_out/phase_1080p_rgba_x16.json
python tools/phase_times.py 1080 1920 3 5 0 1 > gpurun_out/phase_1080p_x1.json 2>&1; cat gpurun_out/phase_1080p_x1.json
python tools/phase_times.py 4320 7680 4 20 1 1 > gpurun_out/phase_8k.json 2>&1; cat gpurun_out/phase_8k.json
