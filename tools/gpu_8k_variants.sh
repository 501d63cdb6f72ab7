#!/bin/bash
# 8K variant sweep: row/column kernel times per compile-time variant (default staged, and unstaged).
mkdir -p gpurun_out
python tools/bench_variants.py --row 0,20 --col 1,21,7 > gpurun_out/variants_8k.jsonl 2>&1; cat gpurun_out/variants_8k.jsonl
echo "--- GBLUR_NO_STAGE=1"
GBLUR_NO_STAGE=1 python tools/bench_variants.py --row 0,20 --col 1 > gpurun_out/variants_8k_nostage.jsonl 2>&1; cat gpurun_out/variants_8k_nostage.jsonl
