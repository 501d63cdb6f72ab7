#!/bin/bash
# 1 GPU: the whole GPU test suite, then the evidence the roofline figures come from — the ncu launch list of the
# bench.py command itself, and --set full captures of the three kernels of the 8K step (Python-free check binary)
mkdir -p gpurun_out/ev
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/ev/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/ev/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/ev/bench_n1.json 2> gpurun_out/ev/bench_n1.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/ev/bench_n1.json").read().strip().splitlines()[-1])
print({k:d.get(k) for k in ("value","ms_per_step","gpu_launches","kernels_ms")}, d["e2e"]["value"], d["e2e"]["mode"], d["roofline"]["frac"], d["secondary"]["band_32k_gray_sigma50"]["ms_per_step"])
PY
# launch list of the same command (cold-cache, serialised: shares, not absolutes)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/ev/launches_bench_8k.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-secondary > gpurun_out/ev/ncu_launches_bench.log 2>&1; echo "ncu launches rc=$?"
tools/ols_gpu_check 4320 7680 20 auto 0 > gpurun_out/ev/plain_ols.log 2>&1 || { cat gpurun_out/ev/plain_ols.log; exit 1; }
cat gpurun_out/ev/plain_ols.log
ncu --set full --clock-control none --import-source on -k regex:ols_row_rgba -s 1 -c 1 -f -o gpurun_out/ev/prof_ols_row \
    tools/ols_gpu_check 4320 7680 20 auto 0 > gpurun_out/ev/ncu_ols_row.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:ols_col_kernel -s 1 -c 1 -f -o gpurun_out/ev/prof_ols_col \
    tools/ols_gpu_check 4320 7680 20 auto 0 > gpurun_out/ev/ncu_ols_col.log 2>&1
ncu --set full --clock-control none -k regex:'altsum_rows_kernel|colsum_kernel|u_rowpass_kernel' -s 6 -c 3 -f -o gpurun_out/ev/prof_ols_pre \
    tools/ols_gpu_check 4320 7680 20 auto 0 > gpurun_out/ev/ncu_ols_pre.log 2>&1
for f in gpurun_out/ev/ncu_ols_*.log; do tail -n 2 $f; done
timeout 300 python tools/step_time.py 2>/dev/null | tee gpurun_out/ev/step_time.json
