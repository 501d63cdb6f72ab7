#!/bin/bash
# One gpurun call: parity tests, bench line, variant sweep, ncu launch list, ncu full capture.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json
python tools/bench_variants.py --row 0,4,5,6 --col 1,7,8,9 > gpurun_out/variants_blocked.jsonl 2>&1
GBLUR_LAYOUT=rowmajor python tools/bench_variants.py --row 0,4,5,6 --col 1,7,8,9 > gpurun_out/variants_rowmajor.jsonl 2>&1
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
python tools/prof_cmd.py 8k 3 > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:line_kernel -s 2 -c 2 -f -o gpurun_out/prof_8k \
    python tools/prof_cmd.py 8k 3 > gpurun_out/ncu_full.log 2>&1
echo done
