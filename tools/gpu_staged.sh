#!/bin/bash
# N-GPU: band-sharded exchange variants of one workload (fused peer stores vs staged copy-engine pushes)
N=${N:-2}
WL=${WL:-band_32k_gray_sigma50}
mkdir -p gpurun_out
if [ -n "$TESTS" ]; then
  timeout 900 python -m pytest tests -m gpu -x -q -k "two_gpu or band_sharded_binding" > gpurun_out/pytest_multi.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_multi.log
fi
for mode in ${MODES:-fused:0 staged:4 staged:8}; do
  ex=${mode%%:*}; k=${mode##*:}
  out=gpurun_out/scale_${WL}_n${N}_${ex}${k}
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
     bench.py --gpus $N --workload $WL --steps 5 --warmup 3 --no-cpu-baseline --exchange $ex --pipeline-chunks $k > $out.json 2> $out.err
  echo "$WL n=$N $ex k=$k rc=$?"; python - <<PY
import json
try:
    d=json.loads(open("$out.json").read().strip().splitlines()[-1])
    print({k:d[k] for k in ("value","ms_per_step","n_gpus")}, d.get("phases_ms"), d["e2e"]["value"])
except Exception as e:
    print("parse failed", e); print(open("$out.err").read()[-2500:])
PY
done
