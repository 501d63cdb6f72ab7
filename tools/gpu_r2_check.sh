#!/bin/bash
# 1 GPU: whole GPU suite, smoke, default bench line, step time
mkdir -p gpurun_out/chk
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/chk/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/chk/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/chk/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/chk/smoke.log
timeout 600 python bench.py > gpurun_out/chk/bench_n1.json 2> gpurun_out/chk/bench_n1.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/chk/bench_n1.json").read().strip().splitlines()[-1])
print({k:d.get(k) for k in ("value","ms_per_step","gpu_launches","kernels_ms")}, d["e2e"]["value"], d["e2e"]["mode"], d["roofline"]["frac"], d["roofline"]["whole_step_frac"])
print(d["secondary"]["band_32k_gray_sigma50"]["ms_per_step"], d["secondary"]["band_32k_gray_sigma50"]["phases_ms"], d.get("parity_check",{}).get("max_abs"))
PY
timeout 300 python tools/step_time.py 2>/dev/null | tee gpurun_out/chk/step_time.json
STEP_ALPHA=0 timeout 300 python tools/step_time.py 2>/dev/null | tee gpurun_out/chk/step_time_alpha0.json
STEP_ALPHA=0 GBLUR_OLS_ALPHA_RESTORE=0 timeout 300 python tools/step_time.py 2>/dev/null | tee gpurun_out/chk/step_time_alpha0_wholeline.json
for shape in "2160 3840 20" "2160 3840 10" "1440 2560 10" "1080 1920 5" "4096 4096 10"; do
  set -- $shape
  STEP_ROWS=$1 STEP_COLS=$2 STEP_SIGMA=$3 timeout 200 python tools/step_time.py 2>/dev/null | tee -a gpurun_out/chk/auto_rule.jsonl
done
