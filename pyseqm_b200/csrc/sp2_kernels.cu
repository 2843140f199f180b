// K3b: SP2 density-matrix purification on packed matrices (pack/unpack folded into the first and last
// kernel).  X0 = (hN I - F)/(hN - h1) from Gershgorin bounds, then X <- X^2 or 2X - X^2 by the trace rule
// until two consecutive |tr X - nocc| < eps; result 2 X.  The multiplications run on the FP64 tensor pipe
// (dgemm_kernels.cu); branch choice, trace and convergence flags stay on the device, the host reads one
// counter per iteration (the reference's `while notconverged.any()`).
// reference: seqm/seqm_functions/SP2.py:3-72, pack.py:7-41
#include "abi_common.cuh"
#include "dgemm.cuh"

namespace seqm {

struct Sp2State {      // per molecule
  double tr, e0, e1, occ;
  int live, niter, square, pad;
};

__device__ __forceinline__ int pk2pad(int q, int nh4) { return q < nh4 ? q : nh4 + 4 * (q - nh4); }

__device__ __forceinline__ double blk_reduce(double v, double* red, int op) {   // op 0 sum, 1 min, 2 max
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double u = __shfl_xor_sync(0xffffffffu, v, o);
    v = op == 0 ? v + u : (op == 1 ? fmin(v, u) : fmax(v, u));
  }
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  double s = red[0];
  for (int q = 1; q < (int)((blockDim.x + 31) >> 5); ++q) s = op == 0 ? s + red[q] : (op == 1 ? fmin(s, red[q]) : fmax(s, red[q]));
  return s;
}

// one CTA per molecule: Gershgorin bounds of the packed matrix, X0, trace
__global__ void __launch_bounds__(256)
sp2_init_kernel(int S, int ld, const double* __restrict__ F, const int32_t* __restrict__ nheavy,
                const int32_t* __restrict__ nhydro, const int32_t* __restrict__ nocc,
                const uint8_t* __restrict__ active, double* __restrict__ X, Sp2State* __restrict__ st,
                int32_t* __restrict__ dims) {
  const int mol = blockIdx.x;
  const int nh4 = 4 * nheavy[mol];
  const int m = nh4 + nhydro[mol];
  if (threadIdx.x == 0) dims[mol] = m;
  const bool on = (!active || active[mol]) && m > 0;
  if (!on) {
    if (threadIdx.x == 0) { st[mol].live = 0; st[mol].niter = 0; }
    return;
  }
  __shared__ double red[8];
  const int n4 = 4 * S;
  const double* Fm = F + (size_t)mol * n4 * n4;
  double* Xm = X + (size_t)mol * ld * ld;
  double lo = 1.0e300, hi = -1.0e300;
  for (int i = threadIdx.x; i < m; i += blockDim.x) {
    const int pi = pk2pad(i, nh4);
    double rad = 0.0, aii = 0.0;
    for (int j = 0; j < m; ++j) {
      const double a = Fm[(size_t)pi * n4 + pk2pad(j, nh4)];
      if (j == i) aii = a; else rad += fabs(a);
    }
    lo = fmin(lo, aii - rad);
    hi = fmax(hi, aii + rad);
  }
  const double h1 = blk_reduce(lo, red, 1);
  const double hN = blk_reduce(hi, red, 2);
  const double sc = hN - h1;
  double tr = 0.0;
  for (int idx = threadIdx.x; idx < ld * ld; idx += blockDim.x) {
    const int i = idx / ld, j = idx - i * ld;
    double x = 0.0;
    if (i < m && j < m) {
      const double a = Fm[(size_t)pk2pad(i, nh4) * n4 + pk2pad(j, nh4)];
      x = ((i == j ? hN : 0.0) - a) / sc;
      if (i == j) tr += x;
    }
    Xm[idx] = x;
  }
  tr = blk_reduce(tr, red, 0);
  if (threadIdx.x == 0) {
    Sp2State s;
    s.occ = (double)nocc[mol];
    s.tr = tr;
    s.e0 = fabs(tr - s.occ);
    s.e1 = s.e0;
    s.live = 1; s.niter = 0; s.square = 0; s.pad = 0;
    st[mol] = s;
  }
}

// after X2 = X X: trace rule, new trace from the diagonals the update will produce, convergence flags
__global__ void __launch_bounds__(256)
sp2_decide_kernel(int ld, double eps, const int32_t* __restrict__ dims, const double* __restrict__ X,
                  const double* __restrict__ X2, Sp2State* __restrict__ st, uint8_t* __restrict__ live,
                  int* __restrict__ nlive) {
  const int mol = blockIdx.x;
  if (!live[mol]) return;
  __shared__ double red[8];
  const int m = dims[mol];
  const double* Xm = X + (size_t)mol * ld * ld;
  const double* Ym = X2 + (size_t)mol * ld * ld;
  double t1 = 0.0, t2 = 0.0;
  for (int i = threadIdx.x; i < m; i += blockDim.x) {
    t1 += Xm[(size_t)i * ld + i];
    t2 += Ym[(size_t)i * ld + i];
  }
  t1 = blk_reduce(t1, red, 0);
  t2 = blk_reduce(t2, red, 0);
  const double occ = st[mol].occ;
  const bool square = fabs(t2 - occ) < fabs(2.0 * t1 - t2 - occ);
  double tn = 0.0;
  for (int i = threadIdx.x; i < m; i += blockDim.x) {
    const double x = Xm[(size_t)i * ld + i], y = Ym[(size_t)i * ld + i];
    tn += square ? y : 2.0 * x - y;
  }
  tn = blk_reduce(tn, red, 0);
  if (threadIdx.x == 0) {
    Sp2State s = st[mol];
    s.square = square ? 1 : 0;
    s.e1 = s.e0;
    s.e0 = fabs(tn - occ);
    s.tr = tn;
    s.niter += 1;
    const bool done = (s.e0 < eps) && (s.e1 < eps);
    s.live = done ? 0 : 1;
    st[mol] = s;
    if (!done) atomicAdd(nlive, 1);
  }
}

// X = X2 or 2X - X2 for the molecules that took this step (live_prev), grid (chunks, nmol)
__global__ void __launch_bounds__(256)
sp2_update_kernel(int ld, const uint8_t* __restrict__ took, const Sp2State* __restrict__ st,
                  double* __restrict__ X, const double* __restrict__ X2) {
  const int mol = blockIdx.y;
  if (!took[mol]) return;
  const bool square = st[mol].square != 0;
  const size_t nn = (size_t)ld * ld;
  double* Xm = X + (size_t)mol * nn;
  const double* Ym = X2 + (size_t)mol * nn;
  for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < nn; idx += (size_t)gridDim.x * blockDim.x) {
    const double y = Ym[idx];
    Xm[idx] = square ? y : 2.0 * Xm[idx] - y;
  }
}

__global__ void sp2_sync_live_kernel(int nmol, const Sp2State* __restrict__ st, uint8_t* __restrict__ live) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b < nmol) live[b] = st[b].live ? 1 : 0;
}

// D = 2 X, unpacked into the padded layout (pack.py:34-41)
__global__ void __launch_bounds__(256)
sp2_finish_kernel(int S, int ld, const double* __restrict__ X, const int32_t* __restrict__ nheavy,
                  const int32_t* __restrict__ dims, const uint8_t* __restrict__ active, const Sp2State* __restrict__ st,
                  double* __restrict__ D, int32_t* __restrict__ niter) {
  const int mol = blockIdx.y;
  if (active && !active[mol]) return;
  const int m = dims[mol], nh4 = 4 * nheavy[mol], n4 = 4 * S;
  if (niter && blockIdx.x == 0 && threadIdx.x == 0) niter[mol] = st[mol].niter;
  const double* Xm = X + (size_t)mol * ld * ld;
  double* Dm = D + (size_t)mol * n4 * n4;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < m * m; idx += gridDim.x * blockDim.x) {
    const int i = idx / m, j = idx - i * m;
    Dm[(size_t)pk2pad(i, nh4) * n4 + pk2pad(j, nh4)] = 2.0 * Xm[(size_t)i * ld + j];
  }
}

static size_t al256(size_t x) { return (x + 255) & ~(size_t)255; }
static int sp2_ld(int mmax) { return (mmax + 3) & ~3; }

}  // namespace seqm

using namespace seqm;

extern "C" {

size_t seqm_sp2_workspace(int64_t nmol, int64_t molsize) {
  const size_t ld = (size_t)4 * molsize;       // upper bound of the packed size
  return 2 * al256((size_t)nmol * ld * ld * 8) + al256(nmol * sizeof(Sp2State)) + al256(nmol * 4) + 2 * al256(nmol) + 256;
}

int seqm_sp2_density(int64_t nmol, int64_t molsize, int mmax, const double* F, const int32_t* nheavy,
                     const int32_t* nhydro, const int32_t* nocc, double eps, const uint8_t* active, double* D,
                     int32_t* niter, void* work, void* stream) {
  SEQM_REQUIRE(F && nheavy && nhydro && nocc && D && work && nmol >= 0 && molsize > 0, "seqm_sp2_density: bad argument");
  if (nmol == 0) return SEQM_OK;
  SEQM_REQUIRE(mmax >= 1 && mmax <= 4 * molsize, "seqm_sp2_density: mmax");
  cudaStream_t st = (cudaStream_t)stream;
  if (eps < 1.0e-7) eps = 1.0e-7;      // SP2.py:21-24 (fp64 branch)
  if (eps > 1.0e-3) eps = 1.0e-3;
  const int B = (int)nmol, S = (int)molsize, ld = sp2_ld(mmax);
  const size_t ldmax = (size_t)4 * molsize;
  char* base = (char*)work;
  double* X = (double*)base; base += al256((size_t)nmol * ldmax * ldmax * 8);
  double* X2 = (double*)base; base += al256((size_t)nmol * ldmax * ldmax * 8);
  Sp2State* state = (Sp2State*)base; base += al256(nmol * sizeof(Sp2State));
  int32_t* dims = (int32_t*)base; base += al256(nmol * 4);
  uint8_t* live = (uint8_t*)base; base += al256(nmol);
  uint8_t* took = (uint8_t*)base; base += al256(nmol);
  int* nlive = (int*)base;
  prof_mark("sp2_init", st);
  sp2_init_kernel<<<B, 256, 0, st>>>(S, ld, F, nheavy, nhydro, nocc, active, X, state, dims);
  SEQM_CHECK_LAUNCH("sp2_init_kernel");
  sp2_sync_live_kernel<<<cdiv(B, 256), 256, 0, st>>>(B, state, live);
  SEQM_CHECK_LAUNCH("sp2_sync_live_kernel");
  const long long stride = (long long)ld * ld;
  int chunks = (int)((stride + 256 * 8 - 1) / (256 * 8));
  if (chunks > 1024) chunks = 1024;
  if (chunks < 1) chunks = 1;
  int* hlive = nullptr;
  SEQM_CUDA(cudaMallocHost(&hlive, sizeof(int)));
  int rc = SEQM_OK;
  for (int it = 0; it < 200; ++it) {
    rc = launch_dgemm_batched(ld, ld, X, X, X2, stride, B, live, dims, st);
    if (rc) break;
    prof_mark("sp2_update+host_sync", st);
    cudaMemsetAsync(nlive, 0, sizeof(int), st);
    cudaMemcpyAsync(took, live, B, cudaMemcpyDeviceToDevice, st);
    sp2_decide_kernel<<<B, 256, 0, st>>>(ld, eps, dims, X, X2, state, live, nlive);
    sp2_update_kernel<<<dim3(chunks, B), 256, 0, st>>>(ld, took, state, X, X2);
    sp2_sync_live_kernel<<<cdiv(B, 256), 256, 0, st>>>(B, state, live);
    cudaMemcpyAsync(hlive, nlive, sizeof(int), cudaMemcpyDeviceToHost, st);
    cudaError_t e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) { set_error(std::string("sp2 loop: ") + cudaGetErrorString(e)); rc = SEQM_ECUDA; break; }
    if (*hlive == 0) break;
  }
  cudaFreeHost(hlive);
  if (rc) return rc;
  int fchunks = (int)(((long long)mmax * mmax + 255) / 256);
  if (fchunks > 512) fchunks = 512;
  prof_mark("sp2_finish", st);
  sp2_finish_kernel<<<dim3(fchunks, B), 256, 0, st>>>(S, ld, X, nheavy, dims, active, state, D, niter);
  SEQM_CHECK_LAUNCH("sp2_finish_kernel");
  prof_mark("end", st);
  return SEQM_OK;
}

}  // extern "C"
