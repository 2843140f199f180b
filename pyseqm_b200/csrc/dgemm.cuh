// internal cross-file declarations of the seqm_b200 library
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

#include "../../include/seqm_b200.h"

namespace seqm {

int launch_dgemm_batched(int n, int ld, const double* A, const double* B, double* C, long long stride, int batch,
                         const uint8_t* active, const int32_t* dims, cudaStream_t st);

int launch_fock(const seqm_topology_t& t, seqm_w_t w, const double* P, const double* M, const double* g_ss,
                const double* g_pp, const double* g_sp, const double* g_p2, const double* h_sp,
                const uint8_t* active, double* F, double* Eelec, double* pairvec, double* scratch, cudaStream_t st,
                const int32_t* nheavy = nullptr, const int32_t* nhydro = nullptr);      // given: energy over orbital rows only
// chunked per-molecule reductions for large molecules (fock_kernels.cu): RED_CHUNK elements per CTA, partial sums in
// scratch [items][chunks], summed in a fixed order
constexpr int RED_CHUNK = 8192;

// (row, column) of element idx of a matrix with ld columns.  A padded matrix of this library has fewer than 2^32
// elements (molsize < 16384: checked where molsize enters the library), so the division is written as a 32-bit one; on
// the size_t index nvcc emits a run-time test for that case around a call into the 64-bit division routine, unrolled
// with the loop (24 call sites in mix_kernel).
__device__ __forceinline__ void row_col(size_t idx, int ld, int& i, int& j) {
  const unsigned u = (unsigned)idx, l = (unsigned)ld;
  const unsigned q = u / l;
  i = (int)q;
  j = (int)(u - q * l);
}
int red_chunks(size_t nn);
int launch_chunk_sum(int items, int chunks, const uint8_t* active, int items_per_mol, const double* part, double* out,
                     const int32_t* out_index, cudaStream_t st);
int launch_elec_energy(int64_t nmol, int64_t molsize, const double* P, const double* F, const double* M,
                       const double* D, int xl, const uint8_t* active, double* E, double* scratch, cudaStream_t st,
                       const int32_t* nheavy = nullptr, const int32_t* nhydro = nullptr);

int launch_eigh_density(int64_t nmol, int64_t molsize, int mmax, const double* F, const int32_t* nheavy,
                        const int32_t* nhydro, const int32_t* nocc, const uint8_t* active, double* P,
                        double* e_out, double* v_out, int32_t* info, double* Vstore, int warm, cudaStream_t st);
size_t eigh_vstore_doubles(int64_t nmol, int mmax);
int launch_eigh_warp(int64_t nmol, int64_t molsize, int mp, const double* F, const int32_t* nheavy, const int32_t* nhydro,
                     const int32_t* nocc, const uint8_t* active, double* P, double* e_out, double* v_out,
                     int32_t* info, double* Vstore, int warm, double tol, cudaStream_t st);


int eigh_cluster_max_m();
int launch_eigh_ring2(int64_t nmol, int64_t molsize, int mp, const double* F, const int32_t* nheavy, const int32_t* nhydro,
                      const int32_t* nocc, const uint8_t* active, double* P, double* e_out, double* v_out, int32_t* info,
                      double* Vstore, int warm, double tol, cudaStream_t st);
size_t eigh_ring2_vstore_doubles(int64_t nmol);
int launch_eigh_cluster(int64_t nmol, int64_t molsize, int mmax, const double* F, const int32_t* nheavy,
                        const int32_t* nhydro, const int32_t* nocc, const uint8_t* active, double* P, double* e_out,
                        double* v_out, int32_t* info, double tol, cudaStream_t st);

double eigh_tolerance();
size_t scf_small_scratch_bytes(int64_t nmol);
int launch_scf_small(const seqm_topology_t& t, seqm_w_t w, const double* M, const double* g_ss, const double* g_pp,
                     const double* g_sp, const double* g_p2, const double* h_sp, const int32_t* nheavy,
                     const int32_t* nhydro, const int32_t* nocc, int mp, int converger, double alpha, double eps,
                     int max_iter, double tol, double* P, double* F, double* E, uint8_t* nc, int32_t* niter,
                     double* pairvec, double* Pold, double* Hpk, double* FOCK, double* FPPF, double* EMAT, double* Vstore,
                     unsigned long long* stat, int32_t* mol_atom0, size_t nn, cudaStream_t st);

}  // namespace seqm
