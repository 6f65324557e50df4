#!/bin/bash
# Tuning builds of the library with extra nvcc flags: scripts/build_variant.sh NAME "-DRT_MIN_BLOCKS=8"
# -> ray_tracer_b200/variants/librt_b200_NAME.so (load with RT_B200_LIB=...)
set -e
cd "$(dirname "$0")/../ray_tracer_b200/csrc"
name=$1; extra=$2
mkdir -p ../variants _build_$name
for f in rt_trace rt_wave rt_bvh rt_api rt_tonemap rt_combine rt_comm rt_order rt_peak; do
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=false -prec-div=true -prec-sqrt=true -ftz=false \
    -Xcompiler -fPIC,-ffp-contract=off -ccbin /usr/bin/g++ $extra -c $f.cu -o _build_$name/$f.o &
done
wait
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../variants/librt_b200_$name.so _build_$name/*.o -cudart static -ldl
rm -rf _build_$name
