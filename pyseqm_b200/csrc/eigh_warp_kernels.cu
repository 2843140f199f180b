// K3a for small molecules (packed size m <= 32): one WARP per molecule, the whole matrix in registers.
// Lane i holds row i of A and row i of the eigenvector matrix V (2 x 32 doubles).  Brent-Luk systolic
// ordering: rotations always act on the adjacent position pairs (2k, 2k+1), so column indices are compile-time
// register indices, and after every round rows and columns are physically permuted (rows by a warp shuffle,
// columns by register renaming) so that all pairs meet in N-1 rounds.  No shared-memory round trips and no
// block barriers inside a sweep; shared memory is only a staging buffer for the load, the warm-start
// transform A <- V0^T A V0 and the density product.
// reference: seqm/seqm_functions/diag.py:102-172 (sym_eig_trunc), pack.py:7-41
#include "abi_common.cuh"

namespace seqm {

constexpr int WN = 32;                 // matrix dimension handled per warp
constexpr int WLD = WN + 1;            // padded smem row
constexpr unsigned FULL = 0xffffffffu;
constexpr int WARP_MAX_SWEEPS = 40;

__host__ __device__ constexpr int bl_src(int j) {
  return j == 0 ? 0 : (j == 2 ? 1 : ((j & 1) == 0 ? j - 2 : (j == WN - 1 ? WN - 2 : j + 2)));
}

__device__ __forceinline__ int pk2pd(int q, int nh4) { return q < nh4 ? q : nh4 + 4 * (q - nh4); }

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
  return v;
}

__global__ void __launch_bounds__(32)
eigh_warp_kernel(int S, const double* __restrict__ F, const int32_t* __restrict__ nheavy,
                 const int32_t* __restrict__ nhydro, const int32_t* __restrict__ nocc,
                 const uint8_t* __restrict__ active, double* __restrict__ P, double* __restrict__ e_out,
                 double* __restrict__ v_out, int32_t* __restrict__ info, double* __restrict__ Vstore, int warm,
                 double tol) {
  const int mol = blockIdx.x;
  if (active && !active[mol]) return;
  const int nh4 = 4 * nheavy[mol];
  const int m = nh4 + nhydro[mol];
  if (m <= 0) return;
  const int n4 = 4 * S;
  const size_t base = (size_t)mol * n4 * n4;
  const int lane = threadIdx.x;
  __shared__ double S0[WN * WLD];
  __shared__ double S1[WN * WLD];
  double a[WN], v[WN];

  // ---- load the packed matrix (upper triangle, like eigh(UPLO='U')): coalesced rows through shared memory
  for (int r = 0; r < WN; ++r) {
    double x = 0.0;
    if (r < m && lane < m) {
      const int rr = r <= lane ? r : lane, cc = r <= lane ? lane : r;
      x = F[base + (size_t)pk2pd(rr, nh4) * n4 + pk2pd(cc, nh4)];
    }
    S0[r * WLD + lane] = x;
  }
  __syncwarp();
#pragma unroll
  for (int j = 0; j < WN; ++j) a[j] = S0[lane * WLD + j];
  __syncwarp();
  double* Vg = Vstore ? Vstore + (size_t)mol * WN * WN : nullptr;
  if (warm && Vg) {
    // V0 -> S0 (coalesced), own row to registers
    for (int r = 0; r < WN; ++r) S0[r * WLD + lane] = Vg[r * WN + lane];
    __syncwarp();
#pragma unroll
    for (int j = 0; j < WN; ++j) v[j] = S0[lane * WLD + j];
    // T = A V0 : row `lane`
    double t[WN];
#pragma unroll
    for (int j = 0; j < WN; ++j) t[j] = 0.0;
#pragma unroll
    for (int k = 0; k < WN; ++k) {
      const double ak = a[k];
#pragma unroll
      for (int j = 0; j < WN; ++j) t[j] = fma(ak, S0[k * WLD + j], t[j]);
    }
#pragma unroll
    for (int j = 0; j < WN; ++j) S1[lane * WLD + j] = t[j];
    __syncwarp();
    // A = V0^T T : row `lane` = sum_k V0[k][lane] T[k][:]
#pragma unroll
    for (int j = 0; j < WN; ++j) a[j] = 0.0;
    for (int k = 0; k < WN; ++k) {
      const double vk = S0[k * WLD + lane];
#pragma unroll
      for (int j = 0; j < WN; ++j) a[j] = fma(vk, S1[k * WLD + j], a[j]);
    }
    __syncwarp();
  } else {
#pragma unroll
    for (int j = 0; j < WN; ++j) v[j] = (j == lane) ? 1.0 : 0.0;
  }

  // source position of this lane's row after the round's permutation
  const int q = bl_src(lane);
  const bool q_even = (q & 1) == 0;
  int sweeps = 0;
  bool last = false;
  for (; sweeps < WARP_MAX_SWEEPS; ++sweeps) {
    double off = 0.0, tot = 0.0;
#pragma unroll
    for (int j = 0; j < WN; ++j) {
      const double x = a[j] * a[j];
      tot += x;
      off += (j == lane) ? 0.0 : x;
    }
    off = warp_sum(off);
    tot = warp_sum(tot);
    if (off == 0.0 || last) break;
    if (off <= tol * tot) last = true;      // quadratic convergence: one more sweep squares the residual
    for (int r = 0; r < WN - 1; ++r) {
      // rotation of this lane's pair (computed on the even lane, shared with the odd one)
      double x = 0.0, y = 0.0;
#pragma unroll
      for (int k = 0; k < WN / 2; ++k)
        if ((lane >> 1) == k) { x = a[2 * k]; y = a[2 * k + 1]; }
      const double yo = __shfl_xor_sync(FULL, y, 1);
      double c = 1.0, s = 0.0;
      if (!(lane & 1) && y != 0.0) {
        // t = sgn(theta)/(|theta| + sqrt(theta^2+1)) with theta = (aqq-app)/(2 apq), written without IEEE
        // division/sqrt: the rotation only has to be orthogonal to rounding (c^2 + s^2 = 1 from one rsqrt)
        const double d = yo - x, b = 2.0 * y;
        const double h2 = d * d + b * b;
        const double hy = h2 * rsqrt(h2);              // sqrt(d^2 + b^2)
        const double den = fabs(d) + hy;
        double inv = rsqrt(den * den);                 // 1/den
        inv = inv * (2.0 - den * inv);                 // one Newton step
        const double tt = ((d >= 0.0) == (b >= 0.0) ? fabs(b) : -fabs(b)) * inv;
        c = rsqrt(tt * tt + 1.0);
        s = tt * c;
      }
      c = __shfl_sync(FULL, c, lane & ~1);
      s = __shfl_sync(FULL, s, lane & ~1);
      // columns: A <- A J, V <- V J
#pragma unroll
      for (int k = 0; k < WN / 2; ++k) {
        const double ck = __shfl_sync(FULL, c, 2 * k), sk = __shfl_sync(FULL, s, 2 * k);
        const double a0 = a[2 * k], a1 = a[2 * k + 1];
        a[2 * k] = ck * a0 - sk * a1;
        a[2 * k + 1] = sk * a0 + ck * a1;
        const double v0 = v[2 * k], v1 = v[2 * k + 1];
        v[2 * k] = ck * v0 - sk * v1;
        v[2 * k + 1] = sk * v0 + ck * v1;
      }
      // rows: A <- J^T A, then rows and columns move to their next-round positions
      const double cq = __shfl_sync(FULL, c, q), sq = __shfl_sync(FULL, s, q);
      double na[WN], nv[WN];
#pragma unroll
      for (int j = 0; j < WN; ++j) {
        const double bj = a[bl_src(j)];
        const double xq = __shfl_sync(FULL, bj, q), xp = __shfl_sync(FULL, bj, q ^ 1);
        na[j] = q_even ? cq * xq - sq * xp : sq * xp + cq * xq;
        nv[j] = v[bl_src(j)];
      }
#pragma unroll
      for (int j = 0; j < WN; ++j) { a[j] = na[j]; v[j] = nv[j]; }
    }
  }
  if (info && lane == 0) info[mol] = sweeps;
  if (Vg) {
#pragma unroll
    for (int j = 0; j < WN; ++j) S0[lane * WLD + j] = v[j];
    __syncwarp();
    for (int r = 0; r < WN; ++r) Vg[r * WN + lane] = S0[r * WLD + lane];
    __syncwarp();
  }

  // eigenvalue of position `lane`; positions that belong to the padding (unit vectors on rows >= m)
  double e = 0.0;
#pragma unroll
  for (int j = 0; j < WN; ++j)
    if (j == lane) e = a[j];
  unsigned dummy = 0u;
#pragma unroll
  for (int p = 0; p < WN; ++p)
    if (__any_sync(FULL, lane >= m && v[p] != 0.0)) dummy |= 1u << p;
  const bool real = !((dummy >> lane) & 1u);
  int rank = 0;
#pragma unroll
  for (int p = 0; p < WN; ++p) {
    const double ep = __shfl_sync(FULL, e, p);
    if (!((dummy >> p) & 1u) && (ep < e || (ep == e && p < lane))) ++rank;
  }
  const int no = nocc[mol];
  const unsigned occmask = __ballot_sync(FULL, real && rank < no);
  if (e_out) {
    if (real) e_out[(size_t)mol * n4 + rank] = e;
    for (int i = m + lane; i < n4; i += 32) e_out[(size_t)mol * n4 + i] = 0.0;
  }
  if (v_out) {
#pragma unroll
    for (int p = 0; p < WN; ++p) {
      const int rp = __shfl_sync(FULL, rank, p);
      if (!((dummy >> p) & 1u) && lane < m) v_out[base + (size_t)lane * n4 + rp] = v[p];
    }
  }
  // P = 2 W W^T with W = occupied columns of V
#pragma unroll
  for (int j = 0; j < WN; ++j) {
    v[j] = ((occmask >> j) & 1u) ? v[j] : 0.0;
    S0[lane * WLD + j] = v[j];
  }
  __syncwarp();
  for (int i2 = 0; i2 < m; ++i2) {
    double acc = 0.0;
#pragma unroll
    for (int j = 0; j < WN; ++j) acc = fma(v[j], S0[i2 * WLD + j], acc);
    S1[i2 * WLD + lane] = 2.0 * acc;          // P[lane][i2] == P[i2][lane]
  }
  __syncwarp();
  for (int i2 = 0; i2 < m; ++i2)
    if (lane < m) P[base + (size_t)pk2pd(i2, nh4) * n4 + pk2pd(lane, nh4)] = S1[i2 * WLD + lane];
}

int launch_eigh_warp(int64_t nmol, int64_t molsize, const double* F, const int32_t* nheavy, const int32_t* nhydro,
                     const int32_t* nocc, const uint8_t* active, double* P, double* e_out, double* v_out,
                     int32_t* info, double* Vstore, int warm, double tol, cudaStream_t st) {
  eigh_warp_kernel<<<(unsigned)nmol, 32, 0, st>>>((int)molsize, F, nheavy, nhydro, nocc, active, P, e_out, v_out, info,
                                                Vstore, warm, tol);
  SEQM_CHECK_LAUNCH("eigh_warp_kernel");
  return SEQM_OK;
}

}  // namespace seqm
