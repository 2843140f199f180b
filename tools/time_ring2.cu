// Cycle accounting of the two-warp ring eigensolver (csrc/eigh_ring2_kernels.cu) on a batch of random symmetric matrices:
// clock64() at the phase boundaries of block 0 while the whole batch runs (cold start, then warm start).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I../pyseqm_b200/csrc -I../include time_ring2.cu
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>
__device__ long long r2_acc[16];
__device__ long long r2_last;
#define SEQM_R2PHASE(i)                                      \
  do {                                                       \
    if (blockIdx.x == 0 && threadIdx.x == 0) {               \
      const long long now_ = clock64();                      \
      if ((i) >= 0) r2_acc[(i)] += now_ - r2_last;           \
      r2_last = now_;                                        \
    }                                                        \
  } while (0)
#include "eigh_ring2_kernels.cu"
namespace seqm {
int g_check_degeneracy = 0, g_eigh_warp_ring = 1;
static thread_local std::string g_err;
void set_error(const std::string& s) { g_err = s; }
std::atomic<unsigned long long> g_launches{0};
}

int main(int argc, char** argv) {
  const int nmol = argc > 1 ? atoi(argv[1]) : 8192;
  const int m = argc > 2 ? atoi(argv[2]) : 48;      // hydrogens only: packed = padded layout, n4 = 4 * S with S = m
  const int S = m, n4 = 4 * S;
  // every "atom" a hydrogen -> packed index q sits at padded index 4 q
  std::vector<double> hF((size_t)nmol * n4 * n4, 0.0);
  unsigned s = 12345;
  for (int b = 0; b < nmol; ++b)
    for (int i = 0; i < m; ++i)
      for (int j = 0; j <= i; ++j) {
        s = s * 1664525u + 1013904223u;
        double x = ((s >> 8) & 0xffff) / 65536.0 - 0.5 + (i == j ? 3.0 * (i % 7) : 0.0);
        hF[(size_t)b * n4 * n4 + (size_t)(4 * i) * n4 + 4 * j] = x;
        hF[(size_t)b * n4 * n4 + (size_t)(4 * j) * n4 + 4 * i] = x;
      }
  double *dF, *dP, *dV;
  int32_t *dnh, *dnhy, *dnocc, *dinfo;
  cudaMalloc(&dF, hF.size() * 8); cudaMalloc(&dP, hF.size() * 8); cudaMalloc(&dV, (size_t)nmol * 64 * 64 * 8);
  cudaMalloc(&dnh, nmol * 4); cudaMalloc(&dnhy, nmol * 4); cudaMalloc(&dnocc, nmol * 4); cudaMalloc(&dinfo, nmol * 4);
  cudaMemcpy(dF, hF.data(), hF.size() * 8, cudaMemcpyHostToDevice);
  std::vector<int32_t> z(nmol, 0), hy(nmol, m), oc(nmol, m / 2);
  cudaMemcpy(dnh, z.data(), nmol * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(dnhy, hy.data(), nmol * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(dnocc, oc.data(), nmol * 4, cudaMemcpyHostToDevice);
  const char* nm[8] = {"load A", "gershgorin", "init G", "sweeps", "rank", "store V/W", "P build", ""};
  for (int warm = 0; warm < 2; ++warm) {
    long long zz[16] = {0};
    cudaMemcpyToSymbol(r2_acc, zz, sizeof(zz));
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a);
    int rc = seqm::launch_eigh_ring2(nmol, S, (m + 1) & ~1, dF, dnh, dnhy, dnocc, nullptr, dP, nullptr, nullptr, dinfo, dV, warm, 1e-16, 0);
    cudaEventRecord(b);
    cudaError_t e = cudaDeviceSynchronize();
    if (rc || e != cudaSuccess) { printf("error %d %s\n", rc, cudaGetErrorString(e)); return 1; }
    float ms; cudaEventElapsedTime(&ms, a, b);
    long long acc[16]; int sw;
    cudaMemcpyFromSymbol(acc, r2_acc, sizeof(acc));
    cudaMemcpy(&sw, dinfo, 4, cudaMemcpyDeviceToHost);
    printf("%s start: %d molecules m=%d: %.3f ms (%.1f ns per eigensolve), block 0: %d sweeps\n", warm ? "warm" : "cold", nmol, m, ms, ms * 1e6 / nmol, sw);
    for (int i = 0; i < 7; ++i) printf("  %-12s %9.1f kcycles\n", nm[i], acc[i] / 1e3);
  }
  return 0;
}
