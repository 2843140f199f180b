// Cycle accounting inside one round of the one-sided ring Jacobi (csrc/eigh_ring.cuh): clock64() at the phase
// boundaries, accumulated over all rounds; NW warps per SM on every SM (argv[1], default 1 block of one warp).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I../pyseqm_b200/csrc -I../include time_ring_round.cu
#include <cstdio>
#include <cstdlib>
#include <string>
__device__ long long tick_acc[8];
__device__ long long tick_last;
__device__ long long g_cnt;
#ifdef NOTICK
#define SEQM_TICK(i)
#else
#define SEQM_TICK(i)                                                    \
  do {                                                                  \
    long long now_ = clock64();                                         \
    if (blockIdx.x == 0 && (threadIdx.x & 31) == 0) {                   \
      if ((i) > 0) tick_acc[(i)] += now_ - tick_last;                   \
      else if (tick_last) tick_acc[0] += now_ - tick_last;              \
      tick_last = now_;                                                 \
    }                                                                   \
  } while (0)
#endif
#include "eigh_ring.cuh"
namespace seqm { int g_check_degeneracy = 0, g_eigh_warp_ring = 1; }

using namespace seqm;
template <int N>
__global__ void __launch_bounds__(32) k(const double* F, double* P, int* sweeps_out) {
  constexpr int LD = WarpLd<N>::value;
  __shared__ __align__(16) double S0[N * LD];
  __shared__ __align__(16) double S1[N * LD];
  const int lane = threadIdx.x;
  for (int r = 0; r < N; ++r)
    if (lane < N) S0[r * LD + lane] = F[r * N + lane];
  __syncwarp();
  long long t0 = clock64();
  int sw = ring_eigh_density<N>(S0, S1, nullptr, 0, 1e-24, N, N / 2, N, nullptr, nullptr);
  long long t1 = clock64();
  if (lane == 0 && blockIdx.x == 0) { *sweeps_out = sw; g_cnt = t1 - t0; }
  if (blockIdx.x == 0)
    for (int r = 0; r < N; ++r)
      if (lane < N) P[r * N + lane] = S1[r * LD + lane];
}

int main(int argc, char** argv) {
  constexpr int N = 30;
  const int nblk = argc > 1 ? atoi(argv[1]) : 1;
  double h[N * N];
  unsigned s = 12345;
  for (int i = 0; i < N; ++i)
    for (int j = 0; j <= i; ++j) {
      s = s * 1664525u + 1013904223u;
      double x = ((s >> 8) & 0xffff) / 65536.0 - 0.5;
      h[i * N + j] = h[j * N + i] = x;
    }
  double *dF, *dP;
  int* dS;
  cudaMalloc(&dF, sizeof(h)); cudaMalloc(&dP, sizeof(h)); cudaMalloc(&dS, 4);
  cudaMemcpy(dF, h, sizeof(h), cudaMemcpyHostToDevice);
  for (int rep = 0; rep < 2; ++rep) {
    long long z[8] = {0}, zl = 0;
    cudaMemcpyToSymbol(tick_acc, z, sizeof(z)); cudaMemcpyToSymbol(tick_last, &zl, 8);
    k<N><<<nblk, 32>>>(dF, dP, dS);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
  }
  long long acc[8], tot; int sw;
  cudaMemcpyFromSymbol(acc, tick_acc, sizeof(acc)); cudaMemcpyFromSymbol(&tot, g_cnt, 8);
  cudaMemcpy(&sw, dS, 4, cudaMemcpyDeviceToHost);
  const double rounds = sw * 31.0;
  printf("blocks %d: sweeps %d rounds %.0f total cycles %lld (%.0f per round incl. everything)\n", nblk, sw, rounds, tot, tot / rounds);
  const char* nm[6] = {"between rounds", "dot + reduce", "tangent", "apply (+ column shuffle)", "norms / scales", "scalar shuffle / exchange"};
  for (int i = 0; i < 6; ++i) printf("  %-28s %8.1f cycles/round\n", nm[i], acc[i] / rounds);
  return 0;
}
