#!/bin/bash
# N-GPU: multi-rank tests + sharded workloads under torchrun
N=${N:-2}
mkdir -p gpurun_out
nvidia-smi -L
timeout 900 python -m pytest tests -m gpu -x -q -k "two_gpu or band_sharded_binding" > gpurun_out/pytest_multi.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_multi.log
for wl in ${WLS:-8k_rgba_sigma20 video_1080p_x512_sigma10 band_16k_gray_sigma50 band_32k_gray_sigma50}; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
     bench.py --gpus $N --workload $wl --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_${wl}_n$N.json 2> gpurun_out/bench_${wl}_n$N.err
  echo "$wl n=$N rc=$?"; cat gpurun_out/bench_${wl}_n$N.json | cut -c1-900; tail -3 gpurun_out/bench_${wl}_n$N.err
done
