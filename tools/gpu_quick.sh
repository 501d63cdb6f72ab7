#!/bin/bash
# Quick GPU check: parity tests, bench line, variant sweep (no profiler).
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json
python tools/bench_variants.py --row ${ROWV:-0,6} --col ${COLV:-1,7,8} > gpurun_out/variants_blocked.jsonl 2>&1
cat gpurun_out/variants_blocked.jsonl
