#!/bin/bash
mkdir -p gpurun_out/q
timeout 600 python -m pytest tests/test_zz_gpu_overlap_save.py -m gpu -x -q > gpurun_out/q/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/q/pytest.log
timeout 600 python bench.py > gpurun_out/q/bench_n1.json 2> gpurun_out/q/bench_n1.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/q/bench_n1.json").read().strip().splitlines()[-1])
print({k:d.get(k) for k in ("value","ms_per_step","gpu_launches","kernels_ms")}, d["e2e"]["value"])
r=d["roofline"]; print({k:r[k] for k in ("kernel","achieved","frac","what","slower_pass_frac","whole_step_frac","traffic")})
PY
