#!/bin/sh
# Instrumented build of the library (per-phase clock64() accounting inside scf_small_kernel, molecule 0):
#   tools/build_timing.sh && SEQM_B200_LIB=tools/_build/libseqm_timing.so python tools/exp_scf_phases.py
set -e
cd "$(dirname "$0")/../pyseqm_b200/csrc"
mkdir -p ../../tools/_build
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo --extended-lambda -Xcompiler -fPIC -DSEQM_SCF_TIMING \
  -shared -o ../../tools/_build/libseqm_timing.so pair_kernels.cu fock_kernels.cu eigh_kernels.cu eigh_warp_kernels.cu eigh_ring2_kernels.cu eigh_cluster_kernels.cu md_kernels.cu parser_kernels.cu \
  dgemm_kernels.cu sp2_kernels.cu scf_kernels.cu scf_small_kernels.cu -lcudart_static -lpthread -ldl -lrt
