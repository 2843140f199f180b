// Cycle accounting inside one Jacobi round of the warp eigensolver (a lone warp): clock64() at the phase
// boundaries of jacobi_round, accumulated over all rounds.   nvcc -arch=sm_100a -O3 -I../pyseqm_b200/csrc -I../include
#include <cstdio>
__device__ long long g_tick[8];
__device__ long long g_cnt;
#define SEQM_TICK(i)                                                    \
  do {                                                                  \
    long long now_ = clock64();                                         \
    if ((threadIdx.x & 31) == 0) {                                      \
      if ((i) > 0) tick_acc[(i)] += now_ - tick_last;                   \
      else if (tick_last) tick_acc[0] += now_ - tick_last;              \
      tick_last = now_;                                                 \
    }                                                                   \
  } while (0)
__device__ long long tick_acc[8];
__device__ long long tick_last;
#include "eigh_warp.cuh"
namespace seqm { int g_check_degeneracy = 0, g_eigh_warp_ring = 0; }

using namespace seqm;
template <int N>
__global__ void k(const double* F, double* P, int* sweeps_out) {
  constexpr int LD = WarpLd<N>::value;
  __shared__ __align__(16) double S0[N * LD];
  __shared__ __align__(16) double S1[N * LD];
  __shared__ __align__(16) double CS[3 * N];
  __shared__ unsigned short TAB[N * 32];
  const int lane = threadIdx.x;
  build_schedule<N>(TAB, lane);
  for (int r = 0; r < N; ++r)
    if (lane < N) S0[r * LD + lane] = F[r * N + lane];
  __syncwarp();
  long long t0 = clock64();
  int sw = warp_eigh_density<N>(S0, S1, CS, TAB, nullptr, 0, 1e-24, N, N / 2, N, nullptr, nullptr);
  long long t1 = clock64();
  if (lane == 0) { *sweeps_out = sw; g_cnt = t1 - t0; }
  for (int r = 0; r < N; ++r)
    if (lane < N) P[r * N + lane] = S1[r * LD + lane];
}

int main() {
  constexpr int N = 30;
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
    k<N><<<1, 32>>>(dF, dP, dS);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
  }
  long long acc[8], tot; int sw;
  cudaMemcpyFromSymbol(acc, tick_acc, sizeof(acc)); cudaMemcpyFromSymbol(&tot, g_cnt, 8);
  cudaMemcpy(&sw, dS, 4, cudaMemcpyDeviceToHost);
  const double rounds = sw * (double)N;
  printf("sweeps %d rounds %.0f total cycles %lld (%.0f per round incl. everything)\n", sw, rounds, tot, tot / rounds);
  const char* nm[7] = {"between rounds", "load+rotation", "CS store+syncwarp", "scale store + row FMAs", "column FMAs", "row publish (STS)", "syncwarp"};
  for (int i = 0; i < 7; ++i) printf("  %-24s %8.1f cycles/round\n", nm[i], acc[i] / rounds);
  return 0;
}
