// K3a-2w stand-alone: sym_eig_trunc + density for packed sizes m <= 32, two warps per molecule (eigh_pair.cuh).
// Used when neither eigenvalues nor eigenvectors are asked for (the SCF loop's density step).
// reference: seqm/seqm_functions/diag.py:102-172 (sym_eig_trunc), pack.py:7-41
#include "eigh_pair.cuh"

namespace seqm {

template <int N>
__global__ void __launch_bounds__(64)
eigh_pair_kernel(int S, const double* __restrict__ F, const int32_t* __restrict__ nheavy,
                 const int32_t* __restrict__ nhydro, const int32_t* __restrict__ nocc,
                 const uint8_t* __restrict__ active, double* __restrict__ P, int32_t* __restrict__ info,
                 double* __restrict__ Vstore, int warm, double tol) {
  constexpr int LD = WarpLd<N>::value;
  const int mol = blockIdx.x;
  if (active && !active[mol]) return;
  const int nh4 = 4 * nheavy[mol];
  const int m = nh4 + nhydro[mol];
  if (m <= 0) return;
  const int n4 = 4 * S;
  const size_t base = (size_t)mol * n4 * n4;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  __shared__ __align__(16) double S0[N * LD];
  __shared__ __align__(16) double S1[N * LD];
  __shared__ __align__(16) double CS[3 * N];
  __shared__ __align__(16) double VB[64];
  __shared__ __align__(16) double red[2];
  __shared__ unsigned short TAB[N * 32];
  if (w == 0) build_schedule<N>(TAB, lane);
  // packed matrix from the upper triangle, like eigh(UPLO='U')
  for (int idx = tid; idx < N * LD; idx += 64) {
    const int r = idx / LD, c = idx - r * LD;
    double x = 0.0;
    if (r < m && c < m) {
      const int rr = r <= c ? r : c, cc = r <= c ? c : r;
      x = F[base + (size_t)pk2pd(rr, nh4) * n4 + pk2pd(cc, nh4)];
    }
    S0[idx] = x;
    S1[idx] = 0.0;
  }
  __syncthreads();
  double* Vg = Vstore ? Vstore + (size_t)mol * WV_LD * WV_LD : nullptr;
  const int sweeps = pair_eigh_density<N>(S0, S1, CS, VB, red, TAB, Vg, warm, tol, m, nocc[mol]);
  if (info && tid == 0) info[mol] = sweeps;
  for (int idx = tid; idx < m * m; idx += 64) {
    const int i = idx / m, j = idx - i * m;
    P[base + (size_t)pk2pd(i, nh4) * n4 + pk2pd(j, nh4)] = S1[i * LD + j];
  }
}

template <int N>
static int launch_pair_n(int64_t nmol, int S, const double* F, const int32_t* nheavy, const int32_t* nhydro,
                         const int32_t* nocc, const uint8_t* active, double* P, int32_t* info, double* Vstore, int warm,
                         double tol, cudaStream_t st) {
  static thread_local bool configured = false;
  if (!configured) {
    SEQM_CUDA(cudaFuncSetAttribute(eigh_pair_kernel<N>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                   (int)cudaSharedmemCarveoutMaxShared));
    configured = true;
  }
  eigh_pair_kernel<N><<<(unsigned)nmol, 64, 0, st>>>(S, F, nheavy, nhydro, nocc, active, P, info, Vstore, warm, tol);
  SEQM_CHECK_LAUNCH("eigh_pair_kernel");
  return SEQM_OK;
}

// mp = even padded size (<= 32); the kernel works on the next multiple of four
int launch_eigh_pair(int64_t nmol, int64_t molsize, int mp, const double* F, const int32_t* nheavy,
                     const int32_t* nhydro, const int32_t* nocc, const uint8_t* active, double* P, int32_t* info,
                     double* Vstore, int warm, double tol, cudaStream_t st) {
#define SEQM_PAIR_CASE(NN) \
  if (mp <= NN) return launch_pair_n<NN>(nmol, (int)molsize, F, nheavy, nhydro, nocc, active, P, info, Vstore, warm, tol, st)
  SEQM_PAIR_CASE(8);
  SEQM_PAIR_CASE(12);
  SEQM_PAIR_CASE(16);
  SEQM_PAIR_CASE(20);
  SEQM_PAIR_CASE(24);
  SEQM_PAIR_CASE(28);
  SEQM_PAIR_CASE(32);
#undef SEQM_PAIR_CASE
  return fail(SEQM_EUNSUPPORTED, "launch_eigh_pair: packed size above 32");
}

}  // namespace seqm
