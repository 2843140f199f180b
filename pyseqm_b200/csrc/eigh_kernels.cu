// K3a: batched symmetric eigensolver + density, one CTA per molecule, everything in shared memory.
//   pack (pack.py:7-21) -> [warm start: A <- V0^T A V0 with the eigenvectors of the previous SCF iteration]
//   -> parallel-order cyclic Jacobi (two-sided, tournament schedule: every round rotates m/2 disjoint index
//   pairs, each thread owns 2x2 tiles so a round needs two barriers) -> rank sort of the eigenvalues ->
//   P = 2 C_occ C_occ^T -> unpack (pack.py:34-41) to the padded layout.
// Works on the PACKED matrix (m = 4 nHeavy + nHydro), so the reference's trick of pushing the padding
// rows above the spectrum (diag.py:116-130) is not needed; orbital energies are returned zero padded like
// diag.py:140-144.  The warm start does not change the result (the decomposition is still converged to
// 1e-14 of the norm); it cuts the sweeps from ~9 to 2-3 because consecutive Fock matrices are close.
// reference: seqm/seqm_functions/diag.py:102-172
#include <mutex>

#include "abi_common.cuh"
#include "dgemm.cuh"

namespace seqm {

constexpr int EIGH_MAX_M = 118;
constexpr int EIGH_WARM_MAX_M = 96;     // three m x m buffers must fit in shared memory for the warm start
constexpr int EIGH_MAX_SWEEPS = 40;

__device__ __forceinline__ int packed_to_padded(int q, int nh4) { return q < nh4 ? q : nh4 + 4 * (q - nh4); }

__device__ __forceinline__ double block_sum(double v, double* red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  double s = 0.0;
  const int nw = (blockDim.x + 31) >> 5;
  for (int q = 0; q < nw; ++q) s += red[q];
  return s;
}

template <int NT, int MINB, bool FAST_T>
__global__ void __launch_bounds__(NT, MINB)
eigh_density_kernel(int S, int ldmax, const double* __restrict__ F, const int32_t* __restrict__ nheavy,
                    const int32_t* __restrict__ nhydro, const int32_t* __restrict__ nocc,
                    const uint8_t* __restrict__ active, double* __restrict__ P, double* __restrict__ e_out,
                    double* __restrict__ v_out, int32_t* __restrict__ info, double* __restrict__ Vstore,
                    int warm, double tol) {
  const int mol = blockIdx.x;
  if (active && !active[mol]) return;
  const int nh4 = 4 * nheavy[mol];
  const int m = nh4 + nhydro[mol];
  const int n4 = 4 * S;
  const size_t base = (size_t)mol * n4 * n4;
  const int tid = threadIdx.x;
  if (m <= 0) return;
  const int mp = (m + 1) & ~1;      // even size for the tournament; the extra index never couples
  const int ld = mp + 1;
  extern __shared__ double sm[];
  const size_t mat = (size_t)ldmax * (ldmax + 1);
  double* A = sm;                   // [mp][ld]
  double* V = A + mat;              // [mp][ld]
  double* T = V + mat;              // [mp][ld] only when Vstore != nullptr
  double* cs = Vstore ? T + mat : T;   // [mp/2][2]
  double* red = cs + ldmax;         // [32]
  int* pq = (int*)(red + 32);       // [mp/2][2]
  int* rank = pq + ldmax;           // [mp]
  int* occ = rank + ldmax;          // [mp]
  double* Vg = Vstore ? Vstore + (size_t)mol * ldmax * ldmax : nullptr;

  for (int idx = tid; idx < mp * mp; idx += NT) {
    const int i = idx / mp, j = idx - i * mp;
    double a = 0.0;
    if (i < m && j < m) {
      const int r = i <= j ? i : j, c = i <= j ? j : i;     // upper triangle, like eigh(UPLO='U')
      a = F[base + (size_t)packed_to_padded(r, nh4) * n4 + packed_to_padded(c, nh4)];
    }
    A[i * ld + j] = a;
    double v = (i == j) ? 1.0 : 0.0;
    if (warm && Vg && i < m && j < m) v = Vg[i * m + j];
    V[i * ld + j] = v;
  }
  __syncthreads();
  if (warm && Vg) {
    // T = A V ; A = V^T T   (upper triangle computed, mirrored)
    for (int idx = tid; idx < m * m; idx += NT) {
      const int i = idx / m, j = idx - i * m;
      double acc = 0.0;
      for (int k = 0; k < m; ++k) acc += A[i * ld + k] * V[k * ld + j];
      T[i * ld + j] = acc;
    }
    __syncthreads();
    for (int idx = tid; idx < m * m; idx += NT) {
      const int i = idx / m, j = idx - i * m;
      if (i > j) continue;
      double acc = 0.0;
      for (int k = 0; k < m; ++k) acc += V[k * ld + i] * T[k * ld + j];
      A[i * ld + j] = acc;
      A[j * ld + i] = acc;
    }
    __syncthreads();
  }

  const int half = mp >> 1;
  int sweeps = 0;
  bool last = false;
  for (; sweeps < EIGH_MAX_SWEEPS; ++sweeps) {
    double off = 0.0, tot = 0.0;
    for (int idx = tid; idx < mp * mp; idx += NT) {
      const int i = idx / mp, j = idx - i * mp;
      const double a = A[i * ld + j];
      tot += a * a;
      if (i != j) off += a * a;
    }
    off = block_sum(off, red);
    tot = block_sum(tot, red);
    // done: exactly diagonal, the announced last sweep has run, or the residual is already at the level a last sweep
    // would be run for anyway (off^2 <= tol^1.5 tot^2, i.e. 1e-12 of the norm at the default tol: quadratic convergence
    // often jumps from 1e-6 straight there, and one more sweep would only polish rounding noise)
    if (off == 0.0 || last || off <= tol * sqrt(tol) * tot) break;
    if (off <= tol * tot) last = true;      // quadratic convergence: one more sweep squares the residual
    for (int r = 0; r < mp - 1; ++r) {
      if (tid < half) {
        int p, q;
        if (tid == 0) { p = mp - 1; q = r; }
        else { p = (r + tid) % (mp - 1); q = (r - tid + mp - 1) % (mp - 1); }
        // (no ordering of p, q: p_k = r + k and q_k = r - k (mod mp - 1) stay consecutive over the pairs k, which is what
        // makes the strided accesses of the update phase below bank-conflict free; the rotation formula is symmetric)
        const double apq = A[p * ld + q];
        double c = 1.0, s = 0.0;
        if (apq != 0.0) {
          // t = sgn(theta) / (|theta| + sqrt(theta^2 + 1)), theta = (aqq - app) / (2 apq), without forming theta
          const double d = A[q * ld + q] - A[p * ld + p], b = 2.0 * apq;
          double tt;
          if (FAST_T) {
            // tangent in FP32 (its 1e-7 relative error only leaves a residual 1e-7 |apq| behind the rotation, which the
            // next sweep removes); c = 1/sqrt(1 + t^2) in FP64 from that same t, so the rotation stays orthogonal to
            // rounding (same scheme as the warp kernel, eigh_warp.cuh)
            const float df = (float)d, bf = (float)b;
            const float h2 = fmaxf(fmaf(df, df, bf * bf), 1.0e-37f);
            const float den = fmaxf(fmaf(h2, __frsqrt_rn(h2), fabsf(df)), 1.0e-37f);
            float tf = __fdividef(df >= 0.0f ? bf : -bf, den);
            tf = fminf(fmaxf(tf, -1.0f), 1.0f);
            tt = (double)tf;
            const double q1 = fma(tt, tt, 1.0);
            double y;
            asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(q1));
            const double e = fma(-q1 * y, y, 1.0);
            c = fma(y * e, fma(0.375, e, 0.5), y);
          } else {
            tt = ((d >= 0.0) == (b >= 0.0) ? fabs(b) : -fabs(b)) / (fabs(d) + sqrt(d * d + b * b));
            c = rsqrt(tt * tt + 1.0);
          }
          s = tt * c;
        }
        cs[2 * tid] = c; cs[2 * tid + 1] = s;
        pq[2 * tid] = p; pq[2 * tid + 1] = q;
      }
      __syncthreads();
      const int nblk = half * half;
      for (int task = tid; task < nblk + mp * half; task += NT) {
        if (task < nblk) {
          // consecutive lanes take consecutive ka: rows pa = r + ka ascend and qa = r - ka descend by one, the row stride
          // ld is odd, so the 16 doubles of a half-warp access fall into distinct bank pairs (ncu r01q: 39 % of the
          // shared-memory wavefronts of the kb-fastest order were conflicts)
          const int kb = task / half, ka = task - kb * half;
          const int2 ra = *reinterpret_cast<const int2*>(pq + 2 * ka), rb = *reinterpret_cast<const int2*>(pq + 2 * kb);
          const double2 ja = *reinterpret_cast<const double2*>(cs + 2 * ka), jb = *reinterpret_cast<const double2*>(cs + 2 * kb);
          const int pa = ra.x, qa = ra.y, pb = rb.x, qb = rb.y;
          const double ca = ja.x, sa = ja.y, cb = jb.x, sb = jb.y;
          const double a00 = A[pa * ld + pb], a01 = A[pa * ld + qb], a10 = A[qa * ld + pb], a11 = A[qa * ld + qb];
          const double b00 = ca * a00 - sa * a10, b01 = ca * a01 - sa * a11;      // rows: J_a^T blk
          const double b10 = sa * a00 + ca * a10, b11 = sa * a01 + ca * a11;
          double c00 = cb * b00 - sb * b01, c01 = sb * b00 + cb * b01;            // columns: blk J_b
          double c10 = cb * b10 - sb * b11, c11 = sb * b10 + cb * b11;
          // exact tangent: the pivot is annihilated by construction; FP32 tangent: keep the (tiny) rotated value, the
          // transformation must stay a similarity transform of what is stored
          if (ka == kb) { if (FAST_T) { c01 = c10 = 0.5 * (c01 + c10); } else { c01 = 0.0; c10 = 0.0; } }
          A[pa * ld + pb] = c00; A[pa * ld + qb] = c01; A[qa * ld + pb] = c10; A[qa * ld + qb] = c11;
        } else {
          const int t2 = task - nblk;
          const int kb = t2 / mp, i = t2 - kb * mp;      // consecutive lanes: consecutive rows of V, same column pair
          const int2 rb = *reinterpret_cast<const int2*>(pq + 2 * kb);
          const double2 jb = *reinterpret_cast<const double2*>(cs + 2 * kb);
          const double v0 = V[i * ld + rb.x], v1 = V[i * ld + rb.y];
          V[i * ld + rb.x] = jb.x * v0 - jb.y * v1;
          V[i * ld + rb.y] = jb.y * v0 + jb.x * v1;
        }
      }
      __syncthreads();
    }
  }
  if (info && tid == 0) info[mol] = sweeps;
  if (Vg)
    for (int idx = tid; idx < m * m; idx += NT) Vg[idx] = V[(idx / m) * ld + (idx % m)];

  // rank of every eigenvalue (ascending, ties by index: a stable sort)
  for (int i = tid; i < m; i += NT) {
    const double ei = A[i * ld + i];
    int rk = 0;
    for (int j = 0; j < m; ++j) {
      const double ej = A[j * ld + j];
      rk += (ej < ei || (ej == ei && j < i)) ? 1 : 0;
    }
    rank[i] = rk;
  }
  __syncthreads();
  const int no = nocc[mol];
  for (int i = tid; i < m; i += NT) {
    if (rank[i] < no) occ[rank[i]] = i;
    if (e_out) e_out[(size_t)mol * n4 + rank[i]] = A[i * ld + i];
  }
  if (e_out)
    for (int i = m + tid; i < n4; i += NT) e_out[(size_t)mol * n4 + i] = 0.0;
  if (v_out) {
    for (int idx = tid; idx < m * m; idx += NT) {
      const int i = idx / m, k = idx - i * m;
      v_out[base + (size_t)i * n4 + rank[k]] = V[i * ld + k];
    }
  }
  __syncthreads();
  // occupied coefficient block C[i][k] into A (the rotated matrix is no longer needed)
  const int nocc_c = no < m ? no : m;
  for (int idx = tid; idx < m * nocc_c; idx += NT) {
    const int i = idx / nocc_c, k = idx - i * nocc_c;
    A[i * ld + k] = V[i * ld + occ[k]];
  }
  __syncthreads();
  for (int idx = tid; idx < m * m; idx += NT) {
    const int i = idx / m, j = idx - i * m;
    if (i > j) continue;
    double acc = 0.0;
    for (int k = 0; k < nocc_c; ++k) acc += A[i * ld + k] * A[j * ld + k];
    acc *= 2.0;
    P[base + (size_t)packed_to_padded(i, nh4) * n4 + packed_to_padded(j, nh4)] = acc;
    P[base + (size_t)packed_to_padded(j, nh4) * n4 + packed_to_padded(i, nh4)] = acc;
  }
}

double g_eigh_tol = 1.0e-16;      // off-diagonal^2 / total^2 at which one last sweep is run (CTA kernel; 1e-24 until r01q: one sweep more, same result to 1e-11 eV)
double g_eigh_tol_warp = 1.0e-16; // same for the warp kernel, whose ordering converges quadratically (tools/exp_eigh_tol.py)
double eigh_tolerance() { return g_eigh_tol_warp; }
int g_eigh_cfg = 3;      // variant hook (seqm_eigh_block_config): bit 0: 256-thread CTAs + FP32 tangent for 32 < m <= 48, bit 1: FP32 tangent for m <= 64
extern int g_eigh_warp_ring;          // eigh_warp_kernels.cu (seqm_eigh_warp_variant)
int g_force_block_jacobi = 0;      // test hook: use the shared-memory kernel also for m <= 32

static size_t eigh_smem_bytes(int mp, bool warm) {
  return ((size_t)(warm ? 3 : 2) * mp * (mp + 1) + mp + 32) * sizeof(double) + (size_t)3 * mp * sizeof(int) + 64;
}

size_t eigh_vstore_doubles(int64_t nmol, int mmax) {
  const int mp = (mmax + 1) & ~1;
  if (mp > 32 && mp <= 64) {      // the two-warp ring kernel keeps [row][64 slots]; the shared-memory kernel mp x mp
    const size_t a = eigh_ring2_vstore_doubles(nmol), b = (size_t)nmol * mp * mp;
    return a > b ? a : b;
  }
  return mmax <= EIGH_WARM_MAX_M ? (size_t)nmol * mp * mp : 0;
}

template <int NT, int MINB, bool FAST_T>
static int launch_cfg(int64_t nmol, size_t smem, cudaStream_t st, int S, int mp, const double* F, const int32_t* nheavy,
                      const int32_t* nhydro, const int32_t* nocc, const uint8_t* active, double* P, double* e_out,
                      double* v_out, int32_t* info, double* Vstore, int warm, double tol) {
  // the attribute belongs to the function, not to the calling thread: it only ever grows (a worker thread of
  // BucketedBatch with a smaller bucket must not lower the limit another thread is about to launch with)
  static std::mutex mtx;
  static size_t configured = 0;
  if (smem > 48 * 1024) {
    std::lock_guard<std::mutex> lock(mtx);
    if (smem > configured) {
      SEQM_CUDA(cudaFuncSetAttribute(eigh_density_kernel<NT, MINB, FAST_T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      configured = smem;
    }
  }
  eigh_density_kernel<NT, MINB, FAST_T><<<(unsigned)nmol, NT, smem, st>>>(S, mp, F, nheavy, nhydro, nocc, active, P, e_out,
                                                                  v_out, info, Vstore, warm, tol);
  SEQM_CHECK_LAUNCH("eigh_density_kernel");
  return SEQM_OK;
}

// Vstore: nullptr (cold start every call) or eigh_vstore_doubles() doubles kept by the caller between calls;
// warm != 0 starts from the eigenvectors stored there by the previous call
int launch_eigh_density(int64_t nmol, int64_t molsize, int mmax, const double* F, const int32_t* nheavy,
                        const int32_t* nhydro, const int32_t* nocc, const uint8_t* active, double* P,
                        double* e_out, double* v_out, int32_t* info, double* Vstore, int warm, cudaStream_t st) {
  if (nmol == 0) return SEQM_OK;
  if (mmax > EIGH_MAX_M) {      // beyond one CTA's shared memory: 8-CTA cluster, one-sided block Jacobi (eigh_cluster_kernels.cu)
    prof_mark("eigh_cluster", st);
    const int rc_ = launch_eigh_cluster(nmol, molsize, mmax, F, nheavy, nhydro, nocc, active, P, e_out, v_out, info, g_eigh_tol_warp, st);
    prof_mark("end", st);
    return rc_;
  }
  if (mmax < 1) mmax = 1;
  if (mmax > EIGH_WARM_MAX_M) { Vstore = nullptr; warm = 0; }
  const int mp = (mmax + 1) & ~1;
  const size_t smem = eigh_smem_bytes(mp, Vstore != nullptr);
  prof_mark("eigh_density", st);
  int rc;
  if (mp <= 32 && !g_force_block_jacobi)
    rc = launch_eigh_warp(nmol, molsize, mp, F, nheavy, nhydro, nocc, active, P, e_out, v_out, info, Vstore, warm, g_eigh_tol_warp, st);
  else if (mp <= 32)
    rc = launch_cfg<256, 7, false>(nmol, smem, st, (int)molsize, mp, F, nheavy, nhydro, nocc, active, P, e_out, v_out, info, Vstore, warm, g_eigh_tol);
  else if (mp <= 64 && g_eigh_warp_ring && !g_force_block_jacobi)
    rc = launch_eigh_ring2(nmol, molsize, mp, F, nheavy, nhydro, nocc, active, P, e_out, v_out, info, Vstore, warm, g_eigh_tol_warp, st);
  else if (mp <= 48 && (g_eigh_cfg & 1))
    rc = launch_cfg<256, 4, true>(nmol, smem, st, (int)molsize, mp, F, nheavy, nhydro, nocc, active, P, e_out, v_out, info, Vstore, warm, g_eigh_tol);
  else if (mp <= 64 && (g_eigh_cfg & 2))
    rc = launch_cfg<512, 2, true>(nmol, smem, st, (int)molsize, mp, F, nheavy, nhydro, nocc, active, P, e_out, v_out, info, Vstore, warm, g_eigh_tol);
  else if (mp <= 64)
    rc = launch_cfg<512, 2, false>(nmol, smem, st, (int)molsize, mp, F, nheavy, nhydro, nocc, active, P, e_out, v_out, info, Vstore, warm, g_eigh_tol);
  else
    rc = launch_cfg<1024, 1, false>(nmol, smem, st, (int)molsize, mp, F, nheavy, nhydro, nocc, active, P, e_out, v_out, info, Vstore, warm, g_eigh_tol);
  prof_mark("end", st);
  return rc;
}

}  // namespace seqm

using namespace seqm;

extern "C" {

int seqm_eigh_max_m(void) { return eigh_cluster_max_m(); }
void seqm_eigh_force_block(int on) { g_force_block_jacobi = on; }
void seqm_eigh_block_config(int bits) { g_eigh_cfg = bits; }

void seqm_eigh_tolerance(double tol) {
  g_eigh_tol = tol > 0.0 ? tol : 1.0e-16;
  g_eigh_tol_warp = tol > 0.0 ? tol : 1.0e-16;
}

int seqm_eigh_density(int64_t nmol, int64_t molsize, int mmax, const double* F, const int32_t* nheavy,
                      const int32_t* nhydro, const int32_t* nocc, const uint8_t* active, double* P, double* e_out,
                      double* v_out, int32_t* info, void* stream) {
  SEQM_REQUIRE(F && nheavy && nhydro && nocc && P && nmol >= 0 && molsize > 0, "seqm_eigh_density: bad argument");
  return launch_eigh_density(nmol, molsize, mmax, F, nheavy, nhydro, nocc, active, P, e_out, v_out, info, nullptr, 0,
                             (cudaStream_t)stream);
}

}  // extern "C"
