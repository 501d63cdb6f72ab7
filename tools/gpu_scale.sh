#!/bin/bash
# scaling runs on an N-GPU box: each workload at the given rank counts
mkdir -p gpurun_out
nvidia-smi -L | wc -l
for wl in ${WLS:-band_32k_gray_sigma50}; do
 for n in ${NS:-8}; do
  if [ "$n" = "1" ]; then
    timeout 900 python bench.py --gpus 1 --workload $wl --steps 5 --warmup 3 --no-cpu-baseline ${EXTRA} > gpurun_out/scale_${wl}_n$n.json 2> gpurun_out/scale_${wl}_n$n.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 \
     bench.py --gpus $n --workload $wl --steps 5 --warmup 3 --no-cpu-baseline ${EXTRA} > gpurun_out/scale_${wl}_n$n.json 2> gpurun_out/scale_${wl}_n$n.err
  fi
  echo "$wl n=$n rc=$?"; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/scale_${wl}_n$n.json").read().strip().splitlines()[-1])
    print({k:d[k] for k in ("value","ms_per_step","n_gpus")}, d.get("phases_ms"), d["e2e"]["value"])
except Exception as e:
    print("parse failed", e); print(open("gpurun_out/scale_${wl}_n$n.err").read()[-1500:])
PY
 done
done
