// (device code shared by eigh_warp_kernels.cu and scf_small_kernels.cu)
// K3a for small molecules (packed size m <= 32): one WARP per molecule.
// Lane i owns row i of A and row i of the eigenvector matrix V in registers (2 x N doubles).  Rows never move
// between lanes; COLUMNS move between register positions.  Ordering: odd-even transposition -- round A pairs
// the positions (0,1),(2,3),..., round B pairs (1,2),(3,4),...; every rotation also swaps its two columns,
// which costs nothing because the rotated values are written to the exchanged registers.  After N rounds the
// column order is reversed and every pair of columns has met exactly once (one sweep); all register indices
// are compile-time constants.  The row half of the similarity transform needs the partner row: rows are
// published in shared memory once per round (double buffered, each row in the slot of the lane that will need it
// next - see jacobi_round) and read back with 128-bit loads; the N/2
// rotation coefficients of a round travel through shared memory as well.  Rotations are applied in the scaled
// (fast Givens) form, one FMA per element.  Per round and lane: 3N FP64 operations + ~40 for the rotation,
// ~1.5N + 4 shared-memory instructions, two warp shuffles (position bookkeeping), two __syncwarp.
// Shared memory per warp: two row buffers N x LD and CS[4N] (rotation coefficients, scales, position table).
// reference: seqm/seqm_functions/diag.py:102-172 (sym_eig_trunc), pack.py:7-41
#pragma once
#include "abi_common.cuh"

#ifndef SEQM_PHASE
#define SEQM_PHASE(i)     // timing builds (tools/build_timing.sh) accumulate clock64() per phase of molecule 0
#endif
#ifndef SEQM_TICK
#define SEQM_TICK(i)      // tools/time_round.cu defines this to sample clock64() inside a round
#endif

namespace seqm {

constexpr unsigned FULL = 0xffffffffu;
constexpr int WARP_MAX_SWEEPS = 40;
constexpr int WV_LD = 32;              // row stride of the eigenvector store kept between SCF iterations
// diag.CHECK_DEGENERACY of the reference (a module-level switch there, a process-wide one here:
// seqm_eigh_check_degeneracy): orbitals degenerate with the HOMO share the remaining electrons equally
extern int g_check_degeneracy;      // host copy, handed to the kernels as an argument
extern int g_eigh_warp_ring;        // process-wide test hook (seqm_eigh_warp_variant): 1 = eigh_ring.cuh, 0 = this file

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

// The schedule of a sweep does not depend on the data: TAB[r][lane] describes round r for the row `lane` when the
// sweep starts from the identity arrangement (column id j at position j):
//   bits 0-4 partner row, bits 5-9 left position L of the pair, bit 10 this row is the left member, bit 11 active.
// A sweep reverses the arrangement; the next sweep is the mirror image (same partners, L -> N-2-L, left <-> right)
// and restores the identity, so one table of N rounds serves all sweeps.
template <int N>
__device__ __forceinline__ void build_schedule(unsigned short* __restrict__ TAB, int lane) {
  const bool on = lane < N;
  int mypos = lane, colid = on ? lane : 0;
#pragma unroll 1
  for (int r = 0; r < N; ++r) {
    const int mp = (r & 1) ? mate_of<N, 1>(mypos) : mate_of<N, 0>(mypos);
    const int ml = (r & 1) ? mate_of<N, 1>(lane) : mate_of<N, 0>(lane);
    const bool act = mp != mypos, left = mypos < mp;
    const int L = (on && act) ? (left ? mypos : mp) : 0;
    const int partner = __shfl_sync(FULL, colid, mp & 31);
    TAB[r * 32 + lane] = (unsigned short)(partner | (L << 5) | ((act && left) ? 1 << 10 : 0) | (act ? 1 << 11 : 0));
    colid = __shfl_sync(FULL, colid, ml & 31);
    mypos = mp;
  }
  __syncwarp();
}

// 1/sqrt(q) for q in [1, 2] without the special-case branches of rsqrt(): hardware seed + one cubic step
__device__ __forceinline__ double rsqrt_unit(double q) {
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(q));
  const double e = fma(-q * y, y, 1.0);
  return fma(y * e, fma(0.375, e, 0.5), y);
}

// One round.  Sr = row buffer holding the current (scaled) A, Sw = buffer that receives the updated rows.
// Rotations are applied in the scaled ("fast Givens") form: the true matrix is D A~ D and the true eigenvector
// matrix V~ D with one scale d per column id, a rotation (c, s = t c) of the columns p, q becomes
//   a~_p <- a~_p - tp a~_q,  a~_q <- a~_q + tq a~_p,   tp = t d_q/d_p, tq = t d_p/d_q,   d_p, d_q <- c d_p, c d_q
// i.e. ONE fused multiply-add per element instead of a multiply and an FMA.  The scales are folded back into
// A~, V~ at the end of every sweep (renormalise()), so they stay in [0.7^N, 1].
// CS[N]: (tp, tq) of the pairs of this round; DSC[id] = (d, 1/d); tab = this round's schedule entry, sigma = 1 in
// the mirrored (odd) sweeps.  The round is branch free.
template <int N, int PAR>
__device__ __forceinline__ void jacobi_round(double (&a)[N], double (&v)[N], const double* __restrict__ Sr,
                                             double* __restrict__ Sw, double* __restrict__ CS,
                                             double2* __restrict__ DSC, unsigned tab, unsigned tab_next, int sigma,
                                             int lane, int rowoff, bool on, double& dmy, double& idmy) {
  constexpr int LD = WarpLd<N>::value;
  constexpr int NPAIR = PAR ? N / 2 - 1 : N / 2;
  SEQM_TICK(0);
  const int partner = tab & 31;
  const bool act = (tab >> 11) & 1;
  const bool left = (((tab >> 10) & 1) != sigma) && act;
  const int L0 = (tab >> 5) & 31;
  const int L = (sigma && on) ? N - 2 - L0 : L0;
  // SLOT SCHEME: in the buffer of round r the row of lane i sits in the slot of its partner of that round, i.e.
  // slot `lane` holds the PARTNER's row.  The 15 LDS.128 of the partner row (the critical path of the round) then
  // read consecutive slots from consecutive lanes - conflict free; the scattered side is the store of the updated
  // row into the slot of the NEXT round's partner, which nothing waits for (ncu r01p: 1.7x excess wavefronts on
  // these loads, short_scoreboard the top stall).
  const int orow = partner * LD;            // own row (two scalars needed)
  const int wrow = (int)(tab_next & 31) * LD;
  const double x0 = Sr[orow + L], x1 = Sr[orow + L + 1];
  const double y0 = Sr[rowoff + L], y1 = Sr[rowoff + L + 1];
  const double2 dpn = DSC[partner];
  double b[N];
#pragma unroll
  for (int jj = 0; jj < N / 2; ++jj) {
    const double2 t = *reinterpret_cast<const double2*>(&Sr[rowoff + 2 * jj]);
    b[2 * jj] = t.x;
    b[2 * jj + 1] = t.y;
  }
  // Rotation from the TRUE 2x2 block (app, apq; apq, aqq).  The tangent t = sgn(d) b / (|d| + sqrt(d^2 + b^2)),
  // d = aqq - app, b = 2 apq, is evaluated in FP32: its error (1e-7 relative) only leaves a residual 1e-7 |apq|
  // behind the rotation, i.e. costs nothing until the off-diagonal norm is at 1e-8 of the matrix norm.  The
  // rotation stays orthogonal to FP64 rounding because c = 1/sqrt(1 + t^2) is computed in FP64 from that same t
  // (needed for the scales only, off the critical path of the round).
  const double dP = left ? dmy : dpn.x, iP = left ? idmy : dpn.y;
  const double dQ = left ? dpn.x : dmy, iQ = left ? dpn.y : idmy;
  const double apq_s = left ? x1 : y1;
  const double app = dP * dP * (left ? x0 : y0), apq = dP * dQ * apq_s, aqq = dQ * dQ * (left ? y1 : x1);
  const double rqp = dQ * iP, rpq = dP * iQ;
  const float df = (float)(aqq - app), bf = (float)(2.0 * apq);
  const float h2 = fmaxf(fmaf(df, df, bf * bf), 1.0e-37f);
  const float den = fmaxf(fmaf(h2, __frsqrt_rn(h2), fabsf(df)), 1.0e-37f);
  float tf = __fdividef(df >= 0.0f ? bf : -bf, den);
  tf = fminf(fmaxf(tf, -1.0f), 1.0f);
  const double t = (act && apq_s != 0.0) ? (double)tf : 0.0;
  const double tp = t * rqp, tq = t * rpq;
  {
    const double q = fma(t, t, 1.0);
    const double c = rsqrt_unit(q);
    dmy *= c;
    idmy *= q * c;
  }
  SEQM_TICK(1);
  if (left) *reinterpret_cast<double2*>(&CS[2 * (L >> 1)]) = make_double2(tp, tq);
  __syncwarp();
  SEQM_TICK(2);
  if (on) DSC[lane] = make_double2(dmy, idmy);
  // rows: row p <- row p - tp row q, row q <- row q + tq row p
  const double sg = left ? -tp : tq;
#pragma unroll
  for (int j = 0; j < N; ++j) a[j] = fma(sg, b[j], a[j]);
  SEQM_TICK(3);
  // columns of A and V, and the two columns of every pair change places
#pragma unroll
  for (int k = 0; k < NPAIR; ++k) {
    const double2 cs = *reinterpret_cast<const double2*>(&CS[2 * k]);
    const int j0 = 2 * k + PAR, j1 = j0 + 1;
    const double a0 = a[j0], a1 = a[j1];
    a[j1] = fma(-cs.x, a1, a0);
    a[j0] = fma(cs.y, a0, a1);
    const double v0 = v[j0], v1 = v[j1];
    v[j1] = fma(-cs.x, v1, v0);
    v[j0] = fma(cs.y, v0, v1);
  }
  SEQM_TICK(4);
  if (on) {
#pragma unroll
    for (int jj = 0; jj < N / 2; ++jj)
      *reinterpret_cast<double2*>(&Sw[wrow + 2 * jj]) = make_double2(a[2 * jj], a[2 * jj + 1]);
  }
  SEQM_TICK(5);
  __syncwarp();
  SEQM_TICK(6);
}

// fold the scales back into A~ and V~ (true values again), reset them to one and publish the rows in S;
// sigma = arrangement after the sweep (1: column id N-1-j at position j)
template <int N>
__device__ __forceinline__ void renormalise(double (&a)[N], double (&v)[N], double* __restrict__ S,
                                            double2* __restrict__ DSC, int sigma, int lane, int slotoff, bool on,
                                            double& dmy, double& idmy) {
#pragma unroll
  for (int j = 0; j < N; ++j) {
    const double dj = DSC[sigma ? N - 1 - j : j].x;
    a[j] *= dmy * dj;
    v[j] *= dj;
  }
  dmy = 1.0;
  idmy = 1.0;
  __syncwarp();
  if (on) {
    DSC[lane] = make_double2(1.0, 1.0);
#pragma unroll
    for (int jj = 0; jj < N / 2; ++jj)
      *reinterpret_cast<double2*>(&S[slotoff + 2 * jj]) = make_double2(a[2 * jj], a[2 * jj + 1]);
  }
  __syncwarp();
}

// Eigen-decomposition of one packed symmetric matrix by one warp, and the density matrix of its lowest `no`
// eigenvectors.
//   in : S0 = the N x LD rows of the matrix (zero beyond m); Vg = this molecule's eigenvector store (or null);
//        warm != 0: start from the eigenvectors kept there by the previous call
//   out: S1 = rows of 2 C_occ C_occ^T (m x m valid); Vg updated; e_out[n4] / v_out[m][n4] (this molecule's,
//        may be null) in the reference's sorted order; returns the number of sweeps
template <int N>
__device__ __forceinline__ int warp_eigh_density(double* __restrict__ S0, double* __restrict__ S1,
                                                 double* __restrict__ CS, const unsigned short* __restrict__ TAB,
                                                 double* __restrict__ Vg, int warm, double tol, int m, int no,
                                                 int n4, double* __restrict__ e_out, double* __restrict__ v_out,
                                                 int check_degeneracy = 0) {
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
    for (int j = 0; j < N; ++j) v[j] = on ? S1[rowoff + j] : 0.0;
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
  } else {
#pragma unroll
    for (int j = 0; j < N; ++j) v[j] = (j == lane) ? 1.0 : 0.0;
  }
  // publish the rows in the slot arrangement of round 0 (slot of the round-0 partner; same in mirrored sweeps)
  const int slot0 = on ? (int)(TAB[lane] & 31) * LD : 0;
  if (on) {
#pragma unroll
    for (int jj = 0; jj < N / 2; ++jj)
      *reinterpret_cast<double2*>(&S0[slot0 + 2 * jj]) = make_double2(a[2 * jj], a[2 * jj + 1]);
  }
  __syncwarp();
  SEQM_PHASE(7);
  // S0 holds the rows of the current A (slot arrangement of round 0)
  double2* DSC = reinterpret_cast<double2*>(CS + N);      // [N] (d, 1/d) per column id
  if (on) DSC[lane] = make_double2(1.0, 1.0);
  double dmy = 1.0, idmy = 1.0;
  __syncwarp();

  int sigma = 0;           // 0: column id j at position j, 1: reversed (after an odd number of sweeps)
  int sweeps = 0;
  bool last = false;
  for (; sweeps < WARP_MAX_SWEEPS; ++sweeps) {
    const int mypos = sigma ? N - 1 - lane : lane;      // position of the diagonal element of this row
    double off = 0.0, tot = 0.0;
#pragma unroll
    for (int j = 0; j < N; ++j) {
      const double x = a[j] * a[j];
      tot += x;
      off += (j == mypos) ? 0.0 : x;
    }
    off = warp_sum(off);
    tot = warp_sum(tot);
    // done: exactly diagonal, the announced last sweep has run, or the residual is already at the level a last sweep
    // would be run for anyway (off^2 <= tol^1.5 tot^2, i.e. 1e-12 of the norm at the default tol: quadratic convergence
    // often jumps from 1e-6 straight there, and one more sweep would only polish rounding noise)
    if (off == 0.0 || last || off <= tol * sqrt(tol) * tot) break;
    if (off <= tol * tot) last = true;      // quadratic convergence: one more sweep squares the residual
    // the schedule entries of the next two rounds are fetched while the current two run (their shared-memory
    // latency sat on the critical path of every round: ncu source view, r01p)
    unsigned t0 = TAB[lane], t1 = TAB[32 + lane];
#pragma unroll 1
    for (int r = 0; r < N; r += 2) {
      const int rn = (r + 2 < N) ? r + 2 : 0;
      const unsigned n0 = TAB[rn * 32 + lane], n1 = TAB[(rn + 1) * 32 + lane];
      jacobi_round<N, 0>(a, v, S0, S1, CS, DSC, t0, t1, sigma, lane, rowoff, on, dmy, idmy);
      jacobi_round<N, 1>(a, v, S1, S0, CS, DSC, t1, n0, sigma, lane, rowoff, on, dmy, idmy);
      t0 = n0;
      t1 = n1;
    }
    sigma ^= 1;
    renormalise<N>(a, v, S0, DSC, sigma, lane, slot0, on, dmy, idmy);
  }
  SEQM_PHASE(8);
  const int colid = on ? (sigma ? N - 1 - lane : lane) : 0;      // column id at position `lane`
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
  const double e = on ? S0[(int)(TAB[colid] & 31) * LD + lane] : 0.0;      // row colid sits in its round-0 partner's slot
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
  // CHECK_DEGENERACY (diag.py:79-98, construct_P): if orbitals beyond the HOMO lie within 1e-14 of it, every orbital of
  // that degenerate set [indx1, indx2) gets the occupation (nocc - indx1) / (indx2 - indx1), the ones below it 1
  double wocc = (real && rank < no) ? 1.0 : 0.0;
  bool smeared = false;
  if (check_degeneracy && no > 0 && no <= m) {
    double ehomo = 0.0;
#pragma unroll
    for (int p = 0; p < N; ++p) {
      const double ep = __shfl_sync(FULL, e, p);
      const int rp = __shfl_sync(FULL, rank, p);
      if (!((dummy >> p) & 1u) && rp == no - 1) ehomo = ep;
    }
    const bool cond = real && fabs(e - ehomo) <= 1.0e-14;
    if (__any_sync(FULL, cond && rank >= no)) {
      int lo = cond ? rank : 0x7fffffff, hi = cond ? rank + 1 : 0;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        lo = min(lo, __shfl_xor_sync(FULL, lo, o));
        hi = max(hi, __shfl_xor_sync(FULL, hi, o));
      }
      wocc = !real ? 0.0 : (rank < lo ? 1.0 : (rank < hi ? (double)(no - lo) / (double)(hi - lo) : 0.0));
      smeared = true;
    }
  }
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
  if (smeared) {      // P = 2 sum_k w_k c_k c_k^T: columns scaled by sqrt(w_k)
    const double sq = sqrt(wocc);
#pragma unroll
    for (int j = 0; j < N; ++j) v[j] *= __shfl_sync(FULL, sq, j);
  } else {
#pragma unroll
    for (int j = 0; j < N; ++j) v[j] = ((occmask >> j) & 1u) ? v[j] : 0.0;
  }
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
  SEQM_PHASE(9);
  return sweeps;
}

}  // namespace seqm
