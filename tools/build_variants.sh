#!/usr/bin/env bash
# Builds experiment variants of the library into build/variants/ (same sources, different compile-time knobs):
#   usage: tools/build_variants.sh name "-DFLAG=..." [name "-D..."]...   then run with RT_B200_LIB=build/variants/lib<name>.so
set -euo pipefail
cd "$(dirname "$0")/.."
make -s all
mkdir -p build/variants
GENCODE="-gencode arch=compute_100a,code=sm_100a"
while [ $# -ge 2 ]; do
  name="$1"; flags="$2"; shift 2
  nvcc $GENCODE -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Iinclude -Icuda_raytracing_b200/csrc/device $flags \
    -c cuda_raytracing_b200/csrc/device/rt_kernels.cu -o build/variants/rt_kernels_$name.o
  nvcc $GENCODE -shared -o build/variants/lib$name.so build/variants/rt_kernels_$name.o build/rt_api.o build/rt_group.o build/rt_pack.o \
    build/rt_world.o build/rt_scenes.o build/rt_host.o -lm -ldl
  echo "built build/variants/lib$name.so ($flags)"
done
