#!/bin/bash
for pair in "12 11" "14 11" "15 11" "18 11" "12 16" "12 19" "15 16"; do set -- $pair; rv=$1; cv=$2
GBLUR_VARIANT_2048=$rv GBLUR_VARIANT_1152=$cv python bench.py --workload video_1080p_x512_sigma10 --steps 3 --warmup 3 --no-cpu-baseline --chunk-frames ${CF:-16} | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('row',$rv,'col',$cv, round(d['value']), round(d['ms_per_step'],2))"
done
