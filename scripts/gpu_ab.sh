# GPU-box A/B helper: the same sweep on tuning builds (ray_tracer_b200/variants/librt_b200_<name>.so, made by
# scripts/build_variant.sh) and the in-tree build ("main"): bash scripts/gpu_ab.sh "base main" ["n_stay,s_min;..."]
mkdir -p gpurun_out/ab
COMBOS=${2:-"8,22"}
for rep in 1 2; do
for v in $1; do
  if [ "$v" = main ]; then unset RT_B200_LIB; else export RT_B200_LIB=$PWD/ray_tracer_b200/variants/librt_b200_$v.so; fi
  for c in "c3_teapot 128" "c4_tree 32" "c5_stress 16" "c2_cube 32"; do
    echo "== $v $c (rep $rep)"; timeout 300 python scripts/sched_sweep.py $c "$COMBOS" 2>&1 | tail -4
  done
done
done 2>&1 | tee gpurun_out/ab/ab.txt
