#!/bin/bash
# A/B of tuning builds on one box: scripts/ab_variants.sh OUTDIR "jobs" name1 name2 ...   (names of ray_tracer_b200/variants/librt_b200_<name>.so)
out=$1; jobs=$2; shift 2
mkdir -p gpurun_out/$out
for round in 1 2; do
  for v in "$@"; do
    echo "== $v (round $round)" | tee -a gpurun_out/$out/ab.log
    RT_B200_LIB=ray_tracer_b200/variants/librt_b200_$v.so python scripts/ab_quick.py $jobs 2>&1 | tee -a gpurun_out/$out/ab.log
  done
done
