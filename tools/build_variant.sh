#!/bin/sh
# Variant build of the library with extra -D flags, for A/B experiments on the GPU box:
#   tools/build_variant.sh name "-DSEQM_PAIRFWD_MINB=3"  ->  tools/_build/libseqm_<name>.so  (use with SEQM_B200_LIB=...)
set -e
name="$1"; shift
cd "$(dirname "$0")/../pyseqm_b200/csrc"
mkdir -p ../../tools/_build
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo --extended-lambda -Xcompiler -fPIC -Xptxas -v "$@" \
  -shared -o ../../tools/_build/libseqm_$name.so pair_kernels.cu fock_kernels.cu eigh_kernels.cu eigh_warp_kernels.cu eigh_pair_kernels.cu \
  dgemm_kernels.cu sp2_kernels.cu scf_kernels.cu scf_small_kernels.cu scf_small2_kernels.cu -lcudart_static -lpthread -ldl -lrt 2> ../../tools/_build/$name.ptxas.log
