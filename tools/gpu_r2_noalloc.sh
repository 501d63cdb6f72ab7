#!/bin/bash
# the no-alloc entry point alone in a fresh process, then the whole GPU suite (it is collected last), then smoke + bench
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -5 gpurun_out/pytest_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; tail -c 600 gpurun_out/bench_n1.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_n1.json").read().strip().splitlines()[-1])
print({k:d.get(k) for k in ("value","ms_per_step","gpu_launches","kernels_ms","parity_check")})
print(d["e2e"]["value"], d["roofline"]["frac"], d["roofline"]["whole_step_frac"], d.get("secondary"))
print(d.get("cpu_baseline"), d.get("clocks"))
PY
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_n1.json 2> gpurun_out/bench_ref_n1.err; echo "ref rc=$?"; cat gpurun_out/bench_ref_n1.json | head -c 900
echo; ls /sys/devices/system/node/ | tr '\n' ' '; nproc
for v in auto 0; do
  if [ "$v" = "0" ]; then export GBLUR_ROW_OLS=0; fi
  timeout 600 python bench.py --workload video_1080p_x512_sigma10 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_video_ols_$v.json 2> gpurun_out/bench_video_ols_$v.err; echo "video ($v) rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_video_ols_$v.json").read().strip().splitlines()[-1])
    print({k:d.get(k) for k in ("value","ms_per_step","gpu_launches")}, d["e2e"]["value"], d.get("phases_ms"), d.get("parity_check"))
except Exception as e:
    print("parse failed", e); print(open("gpurun_out/bench_video_ols_$v.err").read()[-1500:])
PY
done
unset GBLUR_ROW_OLS
