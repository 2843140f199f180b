// (device code of scf_small2_kernels.cu)
// K3a-2w: the warp Jacobi eigensolver of eigh_warp.cuh with TWO warps per molecule.  Thread (w, lane) owns row
// `lane` and the column block [w C, (w+1) C), C = N/2, of A and V in registers, so a round costs every warp half
// the FMAs, half the shared-memory row traffic, and the molecule has two warps in flight (config 2 has only 6.9
// molecules per SM).  Same ordering (odd-even transposition with free column swaps), same scaled rotations, same
// schedule table.  Differences:
//   * the rotation parameters of a round are computed by warp 0 only and travel through CS (as before);
//   * in B rounds the pair (C-1, C) straddles the two blocks: each side recomputes the partner's row-rotated A
//     element from the published rows (one FMA) and reads the partner's V element from VB, which both warps
//     publish at the end of every A round;
//   * __syncwarp becomes a CTA barrier (64 threads), reductions go through shared memory.
// reference: seqm/seqm_functions/diag.py:102-172 (sym_eig_trunc), pack.py:7-41
#pragma once
#include "eigh_warp.cuh"

namespace seqm {

// NOTE: no __restrict__ on the shared-memory pointers of this file: they carry data BETWEEN threads across barriers,
// and with __restrict__ nvcc keeps values loaded before a barrier in registers (seen: a stale reduction result).
// sum over the 64 threads of the CTA; red = 2 doubles of shared memory; contains two barriers
__device__ __forceinline__ double pair_sum(double v, volatile double* red) {
  v = warp_sum(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  return red[0] + red[1];
}

// One round.  a, v: this thread's C columns; Sr/Sw: published (scaled) rows before / after the round;
// CS[N] rotation coefficients (tp, tq) per pair; DSC[id] = (d, 1/d); VB[2][32]: V columns C-1 and C per row.
template <int N, int PAR>
__device__ __forceinline__ void pair_round(double (&a)[N / 2], double (&v)[N / 2], const double* Sr,
                                           double* Sw, double* CS,
                                           double2* DSC, double* VB, unsigned tab, int sigma,
                                           int w, int lane, int rowoff, bool on, double& dmy, double& idmy) {
  constexpr int LD = WarpLd<N>::value;
  constexpr int C = N / 2;
  const int g0 = w * C;
  const int partner = tab & 31;
  const bool act = (tab >> 11) & 1;
  const bool left = (((tab >> 10) & 1) != sigma) && act;
  const int L0 = (tab >> 5) & 31;
  const int L = (sigma && on) ? N - 2 - L0 : L0;
  const int prow = partner * LD;
  if (w == 0) {
    const double x0 = Sr[rowoff + L], x1 = Sr[rowoff + L + 1];
    const double y0 = Sr[prow + L], y1 = Sr[prow + L + 1];
    const double2 dpn = DSC[partner];
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
    const double q = fma(t, t, 1.0);
    const double c = rsqrt_unit(q);
    dmy *= c;
    idmy *= q * c;
    if (left) *reinterpret_cast<double2*>(&CS[2 * (L >> 1)]) = make_double2(t * rqp, t * rpq);
  }
  __syncthreads();
  if (w == 0 && on) DSC[lane] = make_double2(dmy, idmy);
  // rows: row p <- row p - tp row q, row q <- row q + tq row p (this thread's half)
  double sg = 0.0;
  if (act) {
    const double2 own = *reinterpret_cast<const double2*>(&CS[2 * (L >> 1)]);
    sg = left ? -own.x : own.y;
  }
#pragma unroll
  for (int jj = 0; jj < C / 2; ++jj) {
    const double2 b = *reinterpret_cast<const double2*>(&Sr[prow + g0 + 2 * jj]);
    a[2 * jj] = fma(sg, b.x, a[2 * jj]);
    a[2 * jj + 1] = fma(sg, b.y, a[2 * jj + 1]);
  }
  // columns of A and V inside the block; the two columns of a pair change places
  constexpr int NLOC = PAR ? C / 2 - 1 : C / 2;
#pragma unroll
  for (int k = 0; k < NLOC; ++k) {
    const double2 cs = *reinterpret_cast<const double2*>(&CS[2 * (g0 / 2 + k)]);
    const int j0 = 2 * k + PAR, j1 = j0 + 1;
    const double a0 = a[j0], a1 = a[j1];
    a[j1] = fma(-cs.x, a1, a0);
    a[j0] = fma(cs.y, a0, a1);
    const double v0 = v[j0], v1 = v[j1];
    v[j1] = fma(-cs.x, v1, v0);
    v[j0] = fma(cs.y, v0, v1);
  }
  if (PAR == 1) {
    // the pair of positions (C-1, C): p lives in warp 0 (local C-1), q in warp 1 (local 0); after the swap position
    // C-1 holds q' = q + tq p and position C holds p' = p - tp q
    const double2 cs = *reinterpret_cast<const double2*>(&CS[2 * (C / 2 - 1)]);
    if (w == 0) {
      const double aq = on ? fma(sg, Sr[prow + C], Sr[rowoff + C]) : 0.0;       // partner block's element after the row phase
      const double vq = VB[32 + lane];
      a[C - 1] = fma(cs.y, a[C - 1], aq);
      v[C - 1] = fma(cs.y, v[C - 1], vq);
    } else {
      const double ap = on ? fma(sg, Sr[prow + C - 1], Sr[rowoff + C - 1]) : 0.0;
      const double vp = VB[lane];
      a[0] = fma(-cs.x, a[0], ap);
      v[0] = fma(-cs.x, v[0], vp);
    }
  } else {
    // publish the V columns next to the block boundary for the coming B round
    if (w == 0) VB[lane] = v[C - 1];
    else VB[32 + lane] = v[0];
  }
  if (on) {
#pragma unroll
    for (int jj = 0; jj < C / 2; ++jj)
      *reinterpret_cast<double2*>(&Sw[rowoff + g0 + 2 * jj]) = make_double2(a[2 * jj], a[2 * jj + 1]);
  }
  __syncthreads();
}

// fold the scales back into A~ and V~, reset them and publish the rows in S (sigma: arrangement after the sweep)
template <int N>
__device__ __forceinline__ void pair_renormalise(double (&a)[N / 2], double (&v)[N / 2], double* S,
                                                 double2* DSC, int sigma, int w, int lane, int rowoff,
                                                 bool on, double& dmy, double& idmy) {
  constexpr int C = N / 2;
  const int g0 = w * C;
  const double drow = DSC[on ? lane : 0].x;
#pragma unroll
  for (int j = 0; j < C; ++j) {
    const double dj = DSC[sigma ? N - 1 - (g0 + j) : g0 + j].x;
    a[j] *= drow * dj;
    v[j] *= dj;
  }
  dmy = 1.0;
  idmy = 1.0;
  __syncthreads();
  if (on) {
    if (w == 0) DSC[lane] = make_double2(1.0, 1.0);
#pragma unroll
    for (int jj = 0; jj < C / 2; ++jj)
      *reinterpret_cast<double2*>(&S[rowoff + g0 + 2 * jj]) = make_double2(a[2 * jj], a[2 * jj + 1]);
  }
  __syncthreads();
}

// Eigen-decomposition of one packed symmetric matrix by the two warps of a CTA and the density of its lowest `no`
// eigenvectors.  in: S0 = N x LD rows (zero beyond m); Vg: eigenvector store (row stride 32) or null; warm: start
// from it.  out: S1 = rows of 2 C_occ C_occ^T (m x m valid), Vg updated.  Every thread must call it.
// scratch: CS[3N] doubles, VB[64] doubles, red[2] doubles, TAB = schedule (build_schedule<N>)
template <int N>
__device__ __forceinline__ int pair_eigh_density(double* S0, double* S1,
                                                 double* CS, double* VB,
                                                 double* red, const unsigned short* TAB,
                                                 double* Vg, int warm, double tol, int m, int no) {
  constexpr int LD = WarpLd<N>::value;
  constexpr int C = N / 2;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int g0 = w * C;
  const bool on = lane < N;
  const int rowoff = (on ? lane : 0) * LD;
  double a[C], v[C];
#pragma unroll
  for (int jj = 0; jj < C / 2; ++jj) {
    const double2 t = *reinterpret_cast<const double2*>(&S0[rowoff + g0 + 2 * jj]);
    a[2 * jj] = on ? t.x : 0.0;
    a[2 * jj + 1] = on ? t.y : 0.0;
  }
  if (warm && Vg) {
    // V0 -> S1 (coalesced, both warps)
    for (int idx = tid; idx < N * 32; idx += 64) {
      const int r = idx >> 5, cidx = idx & 31;
      if (cidx < N) S1[r * LD + cidx] = Vg[r * WV_LD + cidx];
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < C; ++j) v[j] = on ? S1[rowoff + g0 + j] : 0.0;
    // T = A V0: this thread's C columns of row `lane`
    double t[C];
#pragma unroll
    for (int j = 0; j < C; ++j) t[j] = 0.0;
#pragma unroll 2
    for (int k = 0; k < N; ++k) {
      const double ak = on ? S0[rowoff + k] : 0.0;
#pragma unroll
      for (int jj = 0; jj < C / 2; ++jj) {
        const double2 x = *reinterpret_cast<const double2*>(&S1[k * LD + g0 + 2 * jj]);
        t[2 * jj] = fma(ak, x.x, t[2 * jj]);
        t[2 * jj + 1] = fma(ak, x.y, t[2 * jj + 1]);
      }
    }
    __syncthreads();                       // everybody has read the rows of A
    if (on) {
#pragma unroll
      for (int jj = 0; jj < C / 2; ++jj)
        *reinterpret_cast<double2*>(&S0[rowoff + g0 + 2 * jj]) = make_double2(t[2 * jj], t[2 * jj + 1]);
    }
    __syncthreads();
    // A = V0^T T
#pragma unroll
    for (int j = 0; j < C; ++j) a[j] = 0.0;
#pragma unroll 2
    for (int k = 0; k < N; ++k) {
      const double vk = on ? S1[k * LD + lane] : 0.0;
#pragma unroll
      for (int jj = 0; jj < C / 2; ++jj) {
        const double2 x = *reinterpret_cast<const double2*>(&S0[k * LD + g0 + 2 * jj]);
        a[2 * jj] = fma(vk, x.x, a[2 * jj]);
        a[2 * jj + 1] = fma(vk, x.y, a[2 * jj + 1]);
      }
    }
    __syncthreads();
    if (on) {
#pragma unroll
      for (int jj = 0; jj < C / 2; ++jj)
        *reinterpret_cast<double2*>(&S0[rowoff + g0 + 2 * jj]) = make_double2(a[2 * jj], a[2 * jj + 1]);
    }
  } else {
#pragma unroll
    for (int j = 0; j < C; ++j) v[j] = (g0 + j == lane) ? 1.0 : 0.0;
  }
  double2* DSC = reinterpret_cast<double2*>(CS + N);
  if (w == 0 && on) DSC[lane] = make_double2(1.0, 1.0);
  double dmy = 1.0, idmy = 1.0;
  __syncthreads();

  int sigma = 0, sweeps = 0;
  bool last = false;
  for (; sweeps < WARP_MAX_SWEEPS; ++sweeps) {
    const int mypos = sigma ? N - 1 - lane : lane;
    double off = 0.0, tot = 0.0;
#pragma unroll
    for (int j = 0; j < C; ++j) {
      const double x = a[j] * a[j];
      tot += x;
      off += (g0 + j == mypos) ? 0.0 : x;
    }
    off = pair_sum(off, red);
    tot = pair_sum(tot, red);
    if (off == 0.0 || last) break;
    if (off <= tol * tot) last = true;
#pragma unroll 1
    for (int r = 0; r < N; r += 2) {
      const unsigned t0 = TAB[r * 32 + lane], t1 = TAB[(r + 1) * 32 + lane];
      pair_round<N, 0>(a, v, S0, S1, CS, DSC, VB, t0, sigma, w, lane, rowoff, on, dmy, idmy);
      pair_round<N, 1>(a, v, S1, S0, CS, DSC, VB, t1, sigma, w, lane, rowoff, on, dmy, idmy);
    }
    sigma ^= 1;
    pair_renormalise<N>(a, v, S0, DSC, sigma, w, lane, rowoff, on, dmy, idmy);
  }
  const int colid = on ? (sigma ? N - 1 - lane : lane) : 0;      // column id at position `lane`
  // eigenvalue at POSITION `lane` (both warps compute the full ranking redundantly)
  const double e = on ? S0[colid * LD + lane] : 0.0;
  __syncthreads();                         // S0 rows read; S1 free
  if (Vg) {
    if (on) {
#pragma unroll
      for (int jj = 0; jj < C / 2; ++jj)
        *reinterpret_cast<double2*>(&S1[rowoff + g0 + 2 * jj]) = make_double2(v[2 * jj], v[2 * jj + 1]);
    }
    __syncthreads();
    for (int idx = tid; idx < N * 32; idx += 64) {
      const int r = idx >> 5, cidx = idx & 31;
      if (cidx < N) Vg[r * WV_LD + cidx] = S1[r * LD + cidx];
    }
  }
  // positions that belong to the padding carry unit vectors on rows >= m
  unsigned mine = 0u;
#pragma unroll
  for (int j = 0; j < C; ++j)
    if (__any_sync(FULL, on && lane >= m && v[j] != 0.0)) mine |= 1u << (g0 + j);
  volatile unsigned* ured = reinterpret_cast<volatile unsigned*>(red);
  __syncthreads();
  if (lane == 0) ured[w] = mine;
  __syncthreads();
  const unsigned dummy = ured[0] | ured[1];
  const bool real = on && !((dummy >> lane) & 1u);
  int rank = 0;
#pragma unroll
  for (int p = 0; p < N; ++p) {
    const double ep = __shfl_sync(FULL, e, p);
    if (!((dummy >> p) & 1u) && (ep < e || (ep == e && p < lane))) ++rank;
  }
  const unsigned occmask = __ballot_sync(FULL, real && rank < no);
  // P = 2 W W^T with W = occupied columns of V: W rows -> S0, then row `lane` of P for every second i2
  if (on) {
#pragma unroll
    for (int jj = 0; jj < C / 2; ++jj) {
      const int p0 = g0 + 2 * jj;
      *reinterpret_cast<double2*>(&S0[rowoff + p0]) =
          make_double2(((occmask >> p0) & 1u) ? v[2 * jj] : 0.0, ((occmask >> (p0 + 1)) & 1u) ? v[2 * jj + 1] : 0.0);
    }
  }
  __syncthreads();
  {
    double wr[N];
#pragma unroll
    for (int jj = 0; jj < N / 2; ++jj) {
      const double2 x = *reinterpret_cast<const double2*>(&S0[rowoff + 2 * jj]);
      wr[2 * jj] = x.x;
      wr[2 * jj + 1] = x.y;
    }
    for (int i2 = w; i2 < m; i2 += 2) {
      double acc0 = 0.0, acc1 = 0.0;
#pragma unroll
      for (int jj = 0; jj < N / 2; ++jj) {
        const double2 x = *reinterpret_cast<const double2*>(&S0[i2 * LD + 2 * jj]);
        acc0 = fma(wr[2 * jj], x.x, acc0);
        acc1 = fma(wr[2 * jj + 1], x.y, acc1);
      }
      if (on) S1[i2 * LD + lane] = 2.0 * (acc0 + acc1);
    }
  }
  __syncthreads();
  return sweeps;
}

}  // namespace seqm
