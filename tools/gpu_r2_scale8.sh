#!/bin/bash
# same-lease 1 -> 8 GPU curve of the band-sharded 32768^2 case (default exchange = auto -> halo) and the driver-style
# 8K run at N = 8 (secondary + band_check inside)
mkdir -p gpurun_out/scale
nvidia-smi -L | wc -l
run() {  # n, extra args, out name
  local n=$1 out=$3
  if [ "$n" = "1" ]; then
    timeout 900 python bench.py --gpus 1 $2 > gpurun_out/scale/$out.json 2> gpurun_out/scale/$out.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 \
      bench.py --gpus $n $2 > gpurun_out/scale/$out.json 2> gpurun_out/scale/$out.err
  fi
  echo "$out rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/scale/$out.json").read().strip().splitlines()[-1])
    print({k:d[k] for k in ("value","ms_per_step","n_gpus")}, "e2e", d["e2e"]["value"])
    for k in ("secondary","band_check","parity_check","nvlink"):
        if d.get(k): print(k, json.dumps(d[k])[:700])
    if d.get("phases_ms"): print(d["phases_ms"])
except Exception as e:
    print("parse failed", e); print(open("gpurun_out/scale/$out.err").read()[-1500:])
PY
}
for n in 1 2 4 8; do
  run $n "--workload band_32k_gray_sigma50 --steps 5 --warmup 3 --no-cpu-baseline" band32k_n$n
done
run 8 "--steps 20 --warmup 3" bench8k_n8
run 8 "--workload band_32k_gray_sigma50 --steps 5 --warmup 3 --no-cpu-baseline --exchange staged" band32k_n8_staged
