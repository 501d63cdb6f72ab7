#!/bin/bash
# 8 GPUs: BASELINE.json configs 5 (65536 x 32768 RGB sigma=150, staged exchange) and 3 (512 x 1080p frames) with the final code
mkdir -p gpurun_out/cfg
for wl in band_64kx32k_rgb_sigma150 video_1080p_x512_sigma10; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29514 \
    bench.py --gpus 8 --workload $wl --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/cfg/${wl}_n8.json 2> gpurun_out/cfg/${wl}_n8.err; echo "$wl rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/cfg/${wl}_n8.json").read().strip().splitlines()[-1])
    print({k:d[k] for k in ("value","ms_per_step","n_gpus")}, "e2e", d["e2e"]["value"], d.get("parity_check"), (d.get("phases_ms") or {}).get("exchange"))
except Exception as e:
    print("parse failed", e); print(open("gpurun_out/cfg/${wl}_n8.err").read()[-1500:])
PY
done
