#!/bin/sh
# Variant build of the library for A/B experiments on the GPU box: ONE source recompiled with extra -D flags, linked
# with the objects of the regular build (run `make` in pyseqm_b200/csrc first).
#   tools/build_variant.sh stage pair_kernels.cu -DSEQM_PGX_STAGE=1  ->  tools/_build/libseqm_stage.so
# Use it with SEQM_B200_LIB=tools/_build/libseqm_<name>.so (tools/time_pair_kernels.py takes a list of them).
set -e
name="$1"; src="$2"; shift 2
cd "$(dirname "$0")/../pyseqm_b200/csrc"
mkdir -p ../../tools/_build
obj=../../tools/_build/${src%.cu}_$name.o
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo --extended-lambda -Xcompiler -fPIC -Xptxas -v "$@" \
  -c "$src" -o "$obj" 2> ../../tools/_build/$name.ptxas.log
others=$(ls build/*.o | grep -v "build/${src%.cu}.o")
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../tools/_build/libseqm_$name.so "$obj" $others \
  -lcudart_static -lpthread -ldl -lrt
