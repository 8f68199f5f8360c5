#!/bin/bash
# Developer tool: build libmvstereo with extra flags for aggregate_wave.cu into variants/<name>.so (Dp = 256 kernels only, or
# one other width with e.g. ONLY=WAVE_ONLY_NW=2 in the environment)
#   tools/build_variant.sh <name> "<nvcc -D flags>"       (the other objects come from the last `python -m multiview_b200.build`)
set -e
name=$1; flags=$2
cd "$(dirname "$0")/.."
mkdir -p variants
obj=/tmp/wave_$name.o
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-ffp-contract=off \
   -D${ONLY:-WAVE_ONLY_NW4} $flags -c multiview_b200/csrc/aggregate_wave.cu -o $obj
objs=$(ls multiview_b200/csrc/_obj/*.o | grep -v aggregate_wave.o)
/usr/local/cuda/bin/nvcc -shared -o variants/$name.so $objs $obj -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC
echo built variants/$name.so
