#!/bin/bash
# Tuning aid: build a variant of the library with extra -D flags for ONE source file into build/variants/<name>.so
# (run a probe against it with B200_LIB=build/variants/<name>.so).  usage: build_variant.sh <name> <file.cu> <flags...>
set -e
cd "$(dirname "$0")/../3d-gaussian-splatting_b200/csrc"
name=$1; src=$2; shift 2
out=../../build/variants; mkdir -p $out
make -s >/dev/null
/usr/local/cuda/bin/nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC "$@" -c $src -o $out/$name.o
objs=""; for f in api project_sh binning radix_sort raster_fwd raster_bwd loss adam densify collective; do
  if [ "$f.cu" = "$src" ]; then objs="$objs $out/$name.o"; else objs="$objs $f.o"; fi; done
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $out/$name.so $objs -lcudart
echo built $out/$name.so
