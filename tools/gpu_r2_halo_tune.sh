#!/bin/bash
# halo sharding at N GPUs: timeline of the three streams, and SMs left free for the communication kernels
mkdir -p gpurun_out/halo
N=${N:-8}
for r in ${RESERVES:-0 4 8}; do
  export GBLUR_HALO_RESERVE_SMS=$r
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
    bench.py --gpus $N --workload band_32k_gray_sigma50 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/halo/n${N}_reserve$r.json 2> gpurun_out/halo/n${N}_reserve$r.err
  echo "reserve=$r rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/halo/n${N}_reserve$r.json").read().strip().splitlines()[-1])
    print({k:d[k] for k in ("value","ms_per_step","n_gpus")})
    for i,p in enumerate(d["phases_ms"].get("halo_timeline_ms_per_rank",[])): print(i,p)
except Exception as e:
    print("parse failed", e); print(open("gpurun_out/halo/n${N}_reserve$r.err").read()[-1500:])
PY
done
