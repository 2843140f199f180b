// (device code shared by eigh_warp_kernels.cu and scf_small_kernels.cu)
// K3a for small molecules (packed size m <= 32): one WARP per molecule.
// Lane i owns row i of A and row i of the eigenvector matrix V in registers (2 x N doubles).  Rows never move
// between lanes; COLUMNS move between register positions.  Ordering: odd-even transposition -- round A pairs
// the positions (0,1),(2,3),..., round B pairs (1,2),(3,4),...; every rotation also swaps its two columns,
// which costs nothing because the rotated values are written to the exchanged registers.  After N rounds the
// column order is reversed and every pair of columns has met exactly once (one sweep); all register indices
// are compile-time constants.  The row half of the similarity transform needs the partner row: rows are
// published in shared memory once per round (double buffered) and read back with 128-bit loads; the N/2
// (c,s) pairs of a round travel through shared memory as well.  Per round and lane: 6N FP64 operations,
// ~1.5N + 4 shared-memory instructions, two warp shuffles (position bookkeeping), two __syncwarp.
// reference: seqm/seqm_functions/diag.py:102-172 (sym_eig_trunc), pack.py:7-41
#pragma once
#include "abi_common.cuh"

namespace seqm {

constexpr unsigned FULL = 0xffffffffu;
constexpr int WARP_MAX_SWEEPS = 40;
constexpr int WV_LD = 32;              // row stride of the eigenvector store kept between SCF iterations

__device__ __forceinline__ int pk2pd(int q, int nh4) { return q < nh4 ? q : nh4 + 4 * (q - nh4); }

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
  return v;
}

// row stride (doubles) of the shared row buffers: even (128-bit access) and 2*LD = 4 (mod 32) words so that
// the eight lanes of a quarter warp hit disjoint banks when they read eight consecutive rows
template <int N> struct WarpLd { static constexpr int value = (N <= 18) ? 18 : 34; };

// position paired with position x in a round of parity PAR (x itself when idle)
template <int N, int PAR> __device__ __forceinline__ int mate_of(int x) {
  if (x >= N) return x;
  if (PAR == 0) return x ^ 1;
  return (x == 0 || x == N - 1) ? x : ((x & 1) ? x + 1 : x - 1);
}

// V <- V J for a round of parity PAR whose (c,s) pairs are in CSp (the columns of every pair change places)
template <int N, int PAR>
__device__ __forceinline__ void v_update(double (&v)[N], const double* __restrict__ CSp) {
  constexpr int NPAIR = PAR ? N / 2 - 1 : N / 2;
#pragma unroll
  for (int k = 0; k < NPAIR; ++k) {
    const double2 cs = *reinterpret_cast<const double2*>(&CSp[2 * k]);
    const int j0 = 2 * k + PAR, j1 = j0 + 1;
    const double v0 = v[j0], v1 = v[j1];
    v[j1] = cs.x * v0 - cs.y * v1;
    v[j0] = cs.y * v0 + cs.x * v1;
  }
}

// one round; Sr = row buffer holding the current A, Sw = buffer that receives the updated rows.
// CS = [2][N]: (c,s) pairs of the last A round and of the last B round.  The eigenvector update of the PREVIOUS
// round is executed here, next to the (long, dependent) rotation computation, so that the FP64 pipe has
// independent work while that chain runs.
template <int N, int PAR>
__device__ __forceinline__ void jacobi_round(double (&a)[N], double (&v)[N], const double* __restrict__ Sr,
                                             double* __restrict__ Sw, double* __restrict__ CS, int lane, int rowoff,
                                             bool on, int& mypos, int& colid) {
  constexpr int LD = WarpLd<N>::value;
  constexpr int NPAIR = PAR ? N / 2 - 1 : N / 2;
  double* CSc = CS + PAR * N;
  const double* CSp = CS + (1 - PAR) * N;
  const int mp = mate_of<N, PAR>(mypos);
  const bool act = mp != mypos;
  const bool left = mypos < mp;
  const int L = on ? (left ? mypos : mp) : 0;               // left position of my pair (own position when idle)
  const int partner = __shfl_sync(FULL, colid, mp & 31);    // row that owns the column at the mate position
  const int prow = partner * LD;
  const double x0 = Sr[rowoff + L], x1 = Sr[rowoff + L + 1];
  const double y0 = Sr[prow + L], y1 = Sr[prow + L + 1];
  double b[N];
#pragma unroll
  for (int jj = 0; jj < N / 2; ++jj) {
    const double2 t = *reinterpret_cast<const double2*>(&Sr[prow + 2 * jj]);
    b[2 * jj] = t.x;
    b[2 * jj + 1] = t.y;
  }
  v_update<N, 1 - PAR>(v, CSp);
  // rotation: with r = 1/sqrt(d^2+b^2), u = (1 + |d| r)/2 = cos^2, c = sqrt(u), s = +-|b| r / (2c); branch free,
  // two rsqrt, and c^2 + s^2 = 1 to rounding whatever the accuracy of r
  const double app = left ? x0 : y0, apq = left ? x1 : y1, aqq = left ? y1 : x1;
  const double d = aqq - app, bb = 2.0 * apq;
  const double h2 = fmax(d * d + bb * bb, 1.0e-300);
  const double r = rsqrt(h2);
  const double u = 0.5 + 0.5 * fabs(d) * r;
  const double w = rsqrt(u);
  const bool rot = act && apq != 0.0;
  const double c = rot ? u * w : 1.0;
  const double sa = 0.5 * fabs(bb) * r * w;
  const double s = rot ? (((d >= 0.0) == (bb >= 0.0)) ? sa : -sa) : 0.0;
  if (act && left) *reinterpret_cast<double2*>(&CSc[2 * (L >> 1)]) = make_double2(c, s);
  __syncwarp();
  // rows: A <- J^T A   (row p: c*row_p - s*row_q, row q: s*row_p + c*row_q)
  const double sg = left ? -s : s;
#pragma unroll
  for (int j = 0; j < N; ++j) a[j] = c * a[j] + sg * b[j];
  // columns: A <- A J with the swap
#pragma unroll
  for (int k = 0; k < NPAIR; ++k) {
    const double2 cs = *reinterpret_cast<const double2*>(&CSc[2 * k]);
    const int j0 = 2 * k + PAR, j1 = j0 + 1;
    const double a0 = a[j0], a1 = a[j1];
    a[j1] = cs.x * a0 - cs.y * a1;
    a[j0] = cs.y * a0 + cs.x * a1;
  }
  if (on) {
#pragma unroll
    for (int jj = 0; jj < N / 2; ++jj)
      *reinterpret_cast<double2*>(&Sw[rowoff + 2 * jj]) = make_double2(a[2 * jj], a[2 * jj + 1]);
  }
  colid = __shfl_sync(FULL, colid, mate_of<N, PAR>(lane) & 31);
  mypos = mp;
  __syncwarp();
}

// column index that the first (fake) pending B-round update maps to position j
template <int N> __host__ __device__ constexpr int preswap(int j) {
  return (j == 0 || j >= N - 1) ? j : ((j & 1) ? j + 1 : j - 1);
}

// Eigen-decomposition of one packed symmetric matrix by one warp, and the density matrix of its lowest `no`
// eigenvectors.
//   in : S0 = the N x LD rows of the matrix (zero beyond m); Vg = this molecule's eigenvector store (or null);
//        warm != 0: start from the eigenvectors kept there by the previous call
//   out: S1 = rows of 2 C_occ C_occ^T (m x m valid); Vg updated; e_out[n4] / v_out[m][n4] (this molecule's,
//        may be null) in the reference's sorted order; returns the number of sweeps
template <int N>
__device__ __forceinline__ int warp_eigh_density(double* __restrict__ S0, double* __restrict__ S1,
                                                 double* __restrict__ CS, double* __restrict__ Vg, int warm,
                                                 double tol, int m, int no, int n4, double* __restrict__ e_out,
                                                 double* __restrict__ v_out) {
  constexpr int LD = WarpLd<N>::value;
  const int lane = threadIdx.x & 31;
  const bool on = lane < N;
  const int rowoff = (on ? lane : 0) * LD;
  double a[N], v[N];

#pragma unroll
  for (int jj = 0; jj < N / 2; ++jj) {
    const double2 t = *reinterpret_cast<const double2*>(&S0[rowoff + 2 * jj]);
    a[2 * jj] = on ? t.x : 0.0;
    a[2 * jj + 1] = on ? t.y : 0.0;
  }
  __syncwarp();
  if (warm && Vg) {
    // V0 -> S1 (coalesced), own row to registers
    for (int r = 0; r < N; ++r)
      if (on) S1[r * LD + lane] = Vg[r * WV_LD + lane];
    __syncwarp();
#pragma unroll
    for (int j = 0; j < N; ++j) v[j] = on ? S1[rowoff + preswap<N>(j)] : 0.0;
    // T = A V0 : row `lane`
    double t[N];
#pragma unroll
    for (int j = 0; j < N; ++j) t[j] = 0.0;
#pragma unroll
    for (int k = 0; k < N; ++k) {
      const double ak = a[k];
#pragma unroll
      for (int jj = 0; jj < N / 2; ++jj) {
        const double2 w = *reinterpret_cast<const double2*>(&S1[k * LD + 2 * jj]);
        t[2 * jj] = fma(ak, w.x, t[2 * jj]);
        t[2 * jj + 1] = fma(ak, w.y, t[2 * jj + 1]);
      }
    }
    if (on) {
#pragma unroll
      for (int jj = 0; jj < N / 2; ++jj)
        *reinterpret_cast<double2*>(&S0[rowoff + 2 * jj]) = make_double2(t[2 * jj], t[2 * jj + 1]);
    }
    __syncwarp();
    // A = V0^T T : row `lane` = sum_k V0[k][lane] T[k][:]
#pragma unroll
    for (int j = 0; j < N; ++j) a[j] = 0.0;
#pragma unroll 2
    for (int k = 0; k < N; ++k) {
      const double vk = on ? S1[k * LD + lane] : 0.0;
#pragma unroll
      for (int jj = 0; jj < N / 2; ++jj) {
        const double2 w = *reinterpret_cast<const double2*>(&S0[k * LD + 2 * jj]);
        a[2 * jj] = fma(vk, w.x, a[2 * jj]);
        a[2 * jj + 1] = fma(vk, w.y, a[2 * jj + 1]);
      }
    }
    __syncwarp();
    if (on) {
#pragma unroll
      for (int jj = 0; jj < N / 2; ++jj)
        *reinterpret_cast<double2*>(&S0[rowoff + 2 * jj]) = make_double2(a[2 * jj], a[2 * jj + 1]);
    }
    __syncwarp();
  } else {
#pragma unroll
    for (int j = 0; j < N; ++j) v[j] = (preswap<N>(j) == lane) ? 1.0 : 0.0;
  }
  // S0 holds the rows of the current A; v is stored with the column pairs of a B round exchanged and the
  // pending (deferred) eigenvector update is a B round of identity rotations, which exchanges them back
  if (lane < N / 2) *reinterpret_cast<double2*>(&CS[N + 2 * lane]) = make_double2(1.0, 0.0);
  __syncwarp();

  int mypos = lane;        // register position of column `lane` (the diagonal element of this row)
  int colid = on ? lane : 0;   // column that sits at position `lane`
  int sweeps = 0;
  bool last = false;
  for (; sweeps < WARP_MAX_SWEEPS; ++sweeps) {
    double off = 0.0, tot = 0.0;
#pragma unroll
    for (int j = 0; j < N; ++j) {
      const double x = a[j] * a[j];
      tot += x;
      off += (j == mypos) ? 0.0 : x;
    }
    off = warp_sum(off);
    tot = warp_sum(tot);
    if (off == 0.0 || last) break;
    if (off <= tol * tot) last = true;      // quadratic convergence: one more sweep squares the residual
    for (int r = 0; r < N; r += 2) {
      jacobi_round<N, 0>(a, v, S0, S1, CS, lane, rowoff, on, mypos, colid);
      jacobi_round<N, 1>(a, v, S1, S0, CS, lane, rowoff, on, mypos, colid);
    }
  }
  v_update<N, 1>(v, CS + N);      // flush the deferred update of the last B round
  if (Vg) {
#pragma unroll
    for (int jj = 0; jj < N / 2; ++jj)
      if (on) *reinterpret_cast<double2*>(&S1[rowoff + 2 * jj]) = make_double2(v[2 * jj], v[2 * jj + 1]);
    __syncwarp();
    for (int r = 0; r < N; ++r)
      if (on) Vg[r * WV_LD + lane] = S1[r * LD + lane];
    __syncwarp();
  }

  // eigenvalue at POSITION `lane`: the diagonal element of the row that owns that column
  const double e = on ? S0[colid * LD + lane] : 0.0;
  // positions that belong to the padding carry unit vectors on rows >= m
  unsigned dummy = 0u;
#pragma unroll
  for (int p = 0; p < N; ++p)
    if (__any_sync(FULL, on && lane >= m && v[p] != 0.0)) dummy |= 1u << p;
  const bool real = on && !((dummy >> lane) & 1u);
  int rank = 0;
#pragma unroll
  for (int p = 0; p < N; ++p) {
    const double ep = __shfl_sync(FULL, e, p);
    if (!((dummy >> p) & 1u) && (ep < e || (ep == e && p < lane))) ++rank;
  }
  const unsigned occmask = __ballot_sync(FULL, real && rank < no);
  if (e_out) {
    if (real) e_out[rank] = e;
    for (int i = m + lane; i < n4; i += 32) e_out[i] = 0.0;
  }
  if (v_out) {
#pragma unroll
    for (int p = 0; p < N; ++p) {
      const int rp = __shfl_sync(FULL, rank, p);
      if (!((dummy >> p) & 1u) && lane < m) v_out[(size_t)lane * n4 + rp] = v[p];
    }
  }
  // P = 2 W W^T with W = occupied columns of V
  __syncwarp();
#pragma unroll
  for (int j = 0; j < N; ++j) v[j] = ((occmask >> j) & 1u) ? v[j] : 0.0;
  if (on) {
#pragma unroll
    for (int jj = 0; jj < N / 2; ++jj)
      *reinterpret_cast<double2*>(&S0[rowoff + 2 * jj]) = make_double2(v[2 * jj], v[2 * jj + 1]);
  }
  __syncwarp();
  for (int i2 = 0; i2 < m; ++i2) {
    double acc0 = 0.0, acc1 = 0.0;
#pragma unroll
    for (int jj = 0; jj < N / 2; ++jj) {
      const double2 w = *reinterpret_cast<const double2*>(&S0[i2 * LD + 2 * jj]);
      acc0 = fma(v[2 * jj], w.x, acc0);
      acc1 = fma(v[2 * jj + 1], w.y, acc1);
    }
    if (on) S1[i2 * LD + lane] = 2.0 * (acc0 + acc1);          // P[lane][i2] == P[i2][lane]
  }
  __syncwarp();
  __syncwarp();
  return sweeps;
}

}  // namespace seqm
