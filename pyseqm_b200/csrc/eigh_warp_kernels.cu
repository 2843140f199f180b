// K3a stand-alone: sym_eig_trunc + density for packed sizes m <= 32, one warp per molecule (eigh_warp.cuh).
// reference: seqm/seqm_functions/diag.py:102-172 (sym_eig_trunc), pack.py:7-41
#include "eigh_warp.cuh"
#include "eigh_ring.cuh"

namespace seqm {

int g_check_degeneracy = 0;
int g_eigh_warp_ring = 1;      // 1: one-sided ring Jacobi (eigh_ring.cuh), 0: two-sided odd-even Jacobi (eigh_warp.cuh)

template <int N, bool RING>
__global__ void __launch_bounds__(32)
eigh_warp_kernel(int S, const double* __restrict__ F, const int32_t* __restrict__ nheavy,
                 const int32_t* __restrict__ nhydro, const int32_t* __restrict__ nocc,
                 const uint8_t* __restrict__ active, double* __restrict__ P, double* __restrict__ e_out,
                 double* __restrict__ v_out, int32_t* __restrict__ info, double* __restrict__ Vstore, int warm,
                 double tol, int check_degeneracy) {
  constexpr int LD = WarpLd<N>::value;
  const int mol = blockIdx.x;
  if (active && !active[mol]) return;
  const int nh4 = 4 * nheavy[mol];
  const int m = nh4 + nhydro[mol];
  if (m <= 0) return;
  const int n4 = 4 * S;
  const size_t base = (size_t)mol * n4 * n4;
  const int lane = threadIdx.x;
  const bool on = lane < N;
  __shared__ __align__(16) double S0[N * LD];
  __shared__ __align__(16) double S1[N * LD];
  __shared__ __align__(16) double CS[RING ? 2 : 3 * N];
  __shared__ unsigned short TAB[RING ? 2 : N * 32];
  if constexpr (!RING) build_schedule<N>(TAB, lane);
  // ---- load the packed matrix (upper triangle, like eigh(UPLO='U')) through shared memory
  for (int r = 0; r < N; ++r) {
    double x = 0.0;
    if (r < m && lane < m) {
      const int rr = r <= lane ? r : lane, cc = r <= lane ? lane : r;
      x = F[base + (size_t)pk2pd(rr, nh4) * n4 + pk2pd(cc, nh4)];
    }
    if (on) S0[r * LD + lane] = x;
  }
  __syncwarp();
  __syncwarp();
  double* Vg = Vstore ? Vstore + (size_t)mol * WV_LD * WV_LD : nullptr;
  double* eo = e_out ? e_out + (size_t)mol * n4 : nullptr;
  double* vo = v_out ? v_out + base : nullptr;
  int sweeps;
  if constexpr (RING) sweeps = ring_eigh_density<N>(S0, S1, Vg, warm, tol, m, nocc[mol], n4, eo, vo, check_degeneracy);
  else sweeps = warp_eigh_density<N>(S0, S1, CS, TAB, Vg, warm, tol, m, nocc[mol], n4, eo, vo, check_degeneracy);
  if (info && lane == 0) info[mol] = sweeps;
  for (int i2 = 0; i2 < m; ++i2)
    if (lane < m) P[base + (size_t)pk2pd(i2, nh4) * n4 + pk2pd(lane, nh4)] = S1[i2 * LD + lane];
}

template <int N, bool RING>
static int launch_nv(int64_t nmol, int S, const double* F, const int32_t* nheavy, const int32_t* nhydro,
                    const int32_t* nocc, const uint8_t* active, double* P, double* e_out, double* v_out, int32_t* info,
                    double* Vstore, int warm, double tol, cudaStream_t st) {
  static thread_local bool configured = false;
  if (!configured) {
    SEQM_CUDA(cudaFuncSetAttribute(eigh_warp_kernel<N, RING>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                   (int)cudaSharedmemCarveoutMaxShared));
    configured = true;
  }
  eigh_warp_kernel<N, RING><<<(unsigned)nmol, 32, 0, st>>>(S, F, nheavy, nhydro, nocc, active, P, e_out, v_out, info,
                                                         Vstore, warm, tol, g_check_degeneracy);
  SEQM_CHECK_LAUNCH("eigh_warp_kernel");
  return SEQM_OK;
}

template <int N>
static int launch_n(int64_t nmol, int S, const double* F, const int32_t* nheavy, const int32_t* nhydro,
                    const int32_t* nocc, const uint8_t* active, double* P, double* e_out, double* v_out, int32_t* info,
                    double* Vstore, int warm, double tol, cudaStream_t st) {
  return g_eigh_warp_ring
             ? launch_nv<N, true>(nmol, S, F, nheavy, nhydro, nocc, active, P, e_out, v_out, info, Vstore, warm, tol, st)
             : launch_nv<N, false>(nmol, S, F, nheavy, nhydro, nocc, active, P, e_out, v_out, info, Vstore, warm, tol, st);
}

// mp = even padded size (<= 32)
int launch_eigh_warp(int64_t nmol, int64_t molsize, int mp, const double* F, const int32_t* nheavy,
                     const int32_t* nhydro, const int32_t* nocc, const uint8_t* active, double* P, double* e_out,
                     double* v_out, int32_t* info, double* Vstore, int warm, double tol, cudaStream_t st) {
#define SEQM_WARP_CASE(NN)                                                                                          \
  if (mp <= NN)                                                                                                     \
    return launch_n<NN>(nmol, (int)molsize, F, nheavy, nhydro, nocc, active, P, e_out, v_out, info, Vstore, warm, tol, st)
  SEQM_WARP_CASE(8);
  SEQM_WARP_CASE(12);
  SEQM_WARP_CASE(16);
  SEQM_WARP_CASE(20);
  SEQM_WARP_CASE(24);
  SEQM_WARP_CASE(28);
  SEQM_WARP_CASE(30);
  SEQM_WARP_CASE(32);
#undef SEQM_WARP_CASE
  return fail(SEQM_EUNSUPPORTED, "launch_eigh_warp: packed size above 32");
}

}  // namespace seqm

extern "C" void seqm_eigh_check_degeneracy(int on) {
  seqm::g_check_degeneracy = on ? 1 : 0;
}
extern "C" int seqm_eigh_check_degeneracy_max_m(void) { return 32; }
extern "C" void seqm_eigh_warp_variant(int ring) { seqm::g_eigh_warp_ring = ring ? 1 : 0; }
