#!/bin/bash
# 8 GPUs: the driver-style line (8K per GPU, secondary = 32768^2 halo-sharded with both halo ranges in one launch,
# band_check of both exchanges), then the band workload alone for its e2e and timeline
mkdir -p gpurun_out/final8
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 \
  bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/final8/bench8k_n8.json 2> gpurun_out/final8/bench8k_n8.err; echo "bench n8 rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 \
  bench.py --gpus 8 --workload band_32k_gray_sigma50 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/final8/band32k_n8.json 2> gpurun_out/final8/band32k_n8.err; echo "band n8 rc=$?"
timeout 600 python bench.py --gpus 1 --workload band_32k_gray_sigma50 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/final8/band32k_n1.json 2> gpurun_out/final8/band32k_n1.err; echo "band n1 rc=$?"
python - <<'PY'
import json
for n in ("bench8k_n8","band32k_n8","band32k_n1"):
    try:
        d=json.loads(open(f"gpurun_out/final8/{n}.json").read().strip().splitlines()[-1])
        print(n, {k:d[k] for k in ("value","ms_per_step","n_gpus")}, "e2e", d["e2e"]["value"])
        for k in ("band_check","parity_check"):
            if d.get(k): print(" ", k, json.dumps(d[k])[:400])
        if d.get("secondary"): s=d["secondary"]["band_32k_gray_sigma50"]; print("  secondary", s.get("ms_per_step"), s.get("mpixel_s"), s.get("error"))
        tl=(d.get("phases_ms") or {}).get("halo_timeline_ms_per_rank")
        if tl: print("  rank1", tl[1])
    except Exception as e:
        print(n, "parse failed", e); print(open(f"gpurun_out/final8/{n}.err").read()[-1200:])
PY
