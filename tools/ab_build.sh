#!/bin/bash
# A/B builds of the renderer: tools/ab_build.sh <tag> [extra nvcc flags...]  ->  nerfsafetyvalidation_b200/_lib/ab/libngp_<tag>.so
# (selected at run time with NGP_B200_LIB=...; the .so files are git-ignored and travel with the gpurun snapshot)
set -e
tag=$1; shift
cd "$(dirname "$0")/.."
P=nerfsafetyvalidation_b200
mkdir -p $P/_lib/ab
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --expt-relaxed-constexpr --expt-extended-lambda \
  -Xcompiler -fPIC,-fvisibility=hidden -I include -I $P/csrc "$@" -Xptxas -v -c $P/csrc/fused.cu -o $P/_lib/ab/fused_$tag.o 2> $P/_lib/ab/fused_$tag.log
grep -A2 "k_renderILb0" $P/_lib/ab/fused_$tag.log | tail -2
nvcc -shared -o $P/_lib/ab/libngp_$tag.so $P/_lib/ab/fused_$tag.o $P/_lib/obj/common.o $P/_lib/obj/ops_raymarch.o $P/_lib/obj/ops_encode.o $P/_lib/obj/ops_mlp.o $P/_lib/obj/ops_occupancy.o \
  -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -lcudart
