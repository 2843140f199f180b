// (device code shared by eigh_warp_kernels.cu and scf_small_kernels.cu)
// K3a for small molecules (packed size m <= 32), one WARP per molecule, ONE-SIDED Jacobi in registers.
//
// The two-sided warp kernel (eigh_warp.cuh) publishes every row of A in shared memory once per round and reads its
// partner's row back: ~190 shared-memory wavefronts per round and warp, and with 7 warps per SM the SM's single
// shared-memory pipe was 64 % busy (ncu r02e: 289 M wavefronts in 3.06 M cycles) - the bound of that kernel.  This
// form moves a quarter of those bytes, all of them with warp shuffles, and touches shared memory only before and
// after the iteration:
//   * G = (A + sigma I) V0 (V0 = identity or the eigenvectors kept from the previous SCF iteration), sigma from the
//     Gershgorin bounds so that the spectrum of A + sigma I lies in [w, 2w].  Orthogonalising the COLUMNS of G by plane
//     rotations (Hestenes) diagonalises G^T G = V0^T (A + sigma I)^2 V0; at convergence column k is (lambda_k + sigma)
//     v_k, so the eigenvalue is its norm minus sigma and the eigenvector its direction: V is not accumulated at all.
//   * 2K column slots on K "processors" of R lanes each (R = 2 for m > 16): a processor holds the columns P and Q, a
//     lane H = N / R rows of both (2H doubles).  A rotation needs one dot product (H FMAs + a shuffle across the R
//     lanes; the two norms are carried along, updated with every rotation and recomputed once per
//     sweep) and is applied in the scaled "fast Givens" form as two in-place shears, one FMA per element.
//   * ordering: RECURSIVE RING.  Phase of group size Kg (K, K/2, .., 1): Kg rounds in which P stays and the Q columns
//     of a group of Kg processors move on by one processor (so every P meets every Q of its group); then the two
//     halves of the group exchange one column per processor (all P's to the first half, all Q's to the second) and
//     the phase repeats inside each half.  K + K/2 + .. + 1 = 2K - 1 rounds per sweep, every pair of columns meets
//     exactly once, and only ONE column per processor moves per round: (H + 3) doubles per lane by shuffle, no
//     shared memory, no __syncwarp, all register indices static.  (RingCols::phases also takes group sizes that are
//     not powers of two - used by the two-warp kernel of eigh_ring2_kernels.cu for 20, 24, 28 processors.)
//   * a pair whose cosine is below 1e-13 and whose angle would still be large (|tan| > 1e-3: both the coupling and the
//     norm difference are rounding noise, an exactly degenerate cluster) is not rotated, and the carried norms are
//     updated with the formula that is exact for an inexact tangent.
//   * columns beyond m (and the slots beyond N) are zero vectors: their dot products vanish identically, they are
//     never rotated and are left out of the ranking.
// Stop: a sweep whose rotations met 2 sum gamma^2 <= tol sum alpha^2 was the last one (quadratic convergence: the
// state before it was at sqrt(tol), after it at tol).  The tangent is evaluated in FP32 (as in eigh_warp.cuh); the
// rotation stays orthogonal to FP64 rounding because c is computed in FP64 from that same t.
// reference: seqm/seqm_functions/diag.py:102-172 (sym_eig_trunc, construct_P), pack.py:7-41
#pragma once
#include "eigh_warp.cuh"

namespace seqm {

constexpr double RING_COS_MIN = 1.0e-13;      // |cos| between two columns below which a LARGE rotation is noise ...
constexpr float RING_T_NOISE = 1.0e-3f;       // ... large: |tan| above this

template <int N> struct Ring {
  static constexpr int R = (N <= 8) ? 8 : ((N <= 16) ? 4 : 2);      // lanes per processor
  static constexpr int K = 32 / R;                                  // processors, 2K column slots
  static constexpr int H = N / R;                                   // rows per lane
  // processors that take part in the schedule: all K.  (N / 2 = 10, 12, 14 processors for 20, 24, 28 columns give 2 KA
  // instead of 31 rounds, but their last KA / 2 rounds carry two extra column exchanges each: measured no faster at
  // H <= 14 rows per lane; the two-warp kernel with H >= 20 does use the shorter schedules.)
  static constexpr int KA = K;
  static_assert(N % R == 0 && 2 * K >= N && 2 * KA >= N, "row split / column slots");
};

__device__ __forceinline__ float rsqrt_approx(float x) {
  float y;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <int R> __device__ __forceinline__ double ring_sum(double v) {
#pragma unroll
  for (int o = 1; o < R; o <<= 1) v += __shfl_xor_sync(FULL, v, o);
  return v;
}

// The columns P, Q of one processor as seen by one of its R lanes (H rows of each), their scales (true column =
// d * stored column, fast Givens), their carried squared norms, and the sum of gamma^2 met in the current sweep.
template <int H, int R> struct RingCols {
  double gp[H], gq[H];
  double dp, idp, dq, idq, ap, aq, off;

  // fold the scales into the columns and recompute the norms exactly
  __device__ __forceinline__ void renorm() {
    double sp_ = 0.0, sq_ = 0.0;
#pragma unroll
    for (int i = 0; i < H; ++i) {
      gp[i] *= dp;
      gq[i] *= dq;
      sp_ = fma(gp[i], gp[i], sp_);
      sq_ = fma(gq[i], gq[i], sq_);
    }
    ap = ring_sum<R>(sp_);
    aq = ring_sum<R>(sq_);
    dp = idp = dq = idq = 1.0;
  }

  // one rotation of the columns P, Q; the rotated Q column (and its scalars) then moves on: it is read from lane `from`
  // (the own lane in the round that ends a phase)
  __device__ __forceinline__ void rotate(int from) {
    SEQM_TICK(0);
    double g0 = 0.0, g1 = 0.0, g2_ = 0.0, g3 = 0.0;
#pragma unroll
    for (int i = 0; i < H; ++i) {
      if (H > 16 && (i & 3) == 2) g2_ = fma(gp[i], gq[i], g2_);
      else if (H > 16 && (i & 3) == 3) g3 = fma(gp[i], gq[i], g3);
      else if (i & 1) g1 = fma(gp[i], gq[i], g1);
      else g0 = fma(gp[i], gq[i], g0);
    }
    const double gam = ring_sum<R>((g0 + g1) + (g2_ + g3)) * (dp * dq);
    SEQM_TICK(1);
    // tangent t = sgn(d) b / (|d| + sqrt(d^2 + b^2)), d = beta - alpha, b = 2 gamma, in FP32 with the approximate
    // reciprocal square root / reciprocal (two MUFU instructions; |t| <= 1 up to their 1e-7)
    const float df = (float)(aq - ap), bf = 2.0f * (float)gam;
    const float h2 = fmaxf(fmaf(df, df, bf * bf), 1.0e-37f);
    const float tf = (df >= 0.0f ? bf : -bf) * rcp_approx(fmaf(h2, rsqrt_approx(h2), fabsf(df)));
    // columns already orthogonal to rounding are left alone: inside an exactly degenerate cluster gamma and the norm
    // difference are both noise, the angle would be arbitrary, and such a rotation re-mixes the cluster's couplings to
    // the other columns (the sweep then no longer converges quadratically; tests/dev/proto/onesided_ring.py)
    // (a small angle is harmless and is applied: it is what separates close but distinct eigenvalues)
    const double g2 = gam * gam;
    const bool noise = g2 <= (RING_COS_MIN * RING_COS_MIN) * (ap * aq) && fabsf(tf) > RING_T_NOISE;
    const double t = (gam != 0.0 && !noise) ? (double)tf : 0.0;
    const double tp = t * (dq * idp), tq = t * (dp * idq);
    SEQM_TICK(2);
    // The rotation as two IN-PLACE shears (no temporaries, so no register copies around the 64-bit shuffles):
    //   p <- p - tp q            (the scaled fast-Givens update of column P)
    //   q <- q + tq c^2 p_new    (= (q + tq p_old) / (1 + t^2): the rotated column Q up to the factor 1 + t^2, which goes
    //                             into its scale: d_q <- d_q / c instead of c d_q)
#pragma unroll
    for (int i = 0; i < H; ++i) gp[i] = fma(-tp, gq[i], gp[i]);
    off += g2;
    const double tt = t * t, q = tt + 1.0;
    const double c = rsqrt_unit(q), ic = q * c;
    const double c2 = c * c, tq2 = tq * c2;
    // the moving column leaves element by element as soon as it is rotated: the shuffles overlap the FMAs
#pragma unroll
    for (int i = 0; i < H; ++i) gq[i] = __shfl_sync(FULL, fma(tq2, gp[i], gq[i]), from);
    SEQM_TICK(3);
    // norms after the rotation, exact for ANY t (the shortcut alpha - t gamma holds for the exact tangent only and
    // loses the difference of nearly degenerate columns when t comes from FP32)
    const double tg2 = (t + t) * gam;
    const double ap_n = c2 * (fma(tt, aq, ap) - tg2), aq_n = c2 * (fma(tt, ap, aq) + tg2);
    dp *= c; idp *= ic;
    ap = ap_n;
    SEQM_TICK(4);
    dq = __shfl_sync(FULL, dq * ic, from);
    idq = __shfl_sync(FULL, idq * c, from);
    aq = __shfl_sync(FULL, aq_n, from);
    SEQM_TICK(5);
  }

  // two processors exchange one column inside the warp: the `first` one gives its Q and takes the partner's P into its
  // Q slot, the other gives its P and takes the partner's Q into its P slot; plane = the partner's lane with the same
  // row part (the own lane: nothing changes)
  __device__ __forceinline__ void exchange(bool first, int plane) {
#pragma unroll
    for (int i = 0; i < H; ++i) {
      const double got = __shfl_sync(FULL, first ? gq[i] : gp[i], plane);
      gq[i] = first ? got : gq[i];
      gp[i] = first ? gp[i] : got;
    }
    const double g_d = __shfl_sync(FULL, first ? dq : dp, plane);
    const double g_i = __shfl_sync(FULL, first ? idq : idp, plane);
    const double g_a = __shfl_sync(FULL, first ? aq : ap, plane);
    dq = first ? g_d : dq; idq = first ? g_i : idq; aq = first ? g_a : aq;
    dp = first ? dp : g_d; idp = first ? idp : g_i; ap = first ? ap : g_a;
  }

  // g rounds of a ring of the g processors base .. base + g - 1 (P stays, Q moves on by one processor after every round
  // but the last); idle: a processor outside the schedule (it holds empty columns and shuffles with itself)
  __device__ __forceinline__ void ring(int g, int base, int kw, int h, int lane, bool idle) {
    const int src = idle ? lane : (base + (kw - base + g - 1) % g) * R + h;      // previous processor of the group, same row part
#pragma unroll 1
    for (int r = 0; r < g; ++r) rotate(r + 1 < g ? src : lane);
  }

  // All pairs among the 2 K0 columns held by the processors 0 .. K0 - 1 of one warp (kw = processor index inside the
  // warp; processors >= K0 idle).  Group of g processors: ring(g) lets every P meet every Q of the group; g even: the
  // halves exchange (P's to the first half, Q's to the second) and each half is a group of g / 2; g odd (3, 5, 7 - the
  // end of K0 = 12, 10, 14): round robin of the g P's and the g Q's in g more rounds - in round r processor i swaps one
  // column with processor (2r - i) mod g, rotates, and swaps back (processor r rotates its own pair once more).
  // Rounds: 2 K0 - 1 for K0 = 2^k, 2 K0 for K0 = 10, 12, 14 (tests/dev/proto/ring_schedule.py checks the coverage).
  // ONE copy of the round in the instruction stream (the one-launch SCF kernel around it is large and its warps run
  // different phases: unrolled phases showed as no_instruction stalls).
  __device__ __forceinline__ void phases(int K0, int kw, int h, int lane) {
    const bool idle = kw >= K0;
    int g = K0, base = 0;
#pragma unroll 1
    for (;;) {
      ring(g, base, kw, h, lane, idle);
      if (g == 1) break;
      if ((g & 1) == 0) {
        const int half = g >> 1;
        const bool first = kw - base < half;
        exchange(first, idle ? lane : (first ? kw + half : kw - half) * R + h);
        if (!first) base += half;
        g = half;
      } else {
        const int hh = (g - 1) >> 1, i = kw - base;
#pragma unroll 1
        for (int r = 0; r < g; ++r) {
          const int d = (i - r + g) % g;      // 0: the bye
          const bool first = d > hh;
          const int plane = (idle || d == 0) ? lane : (base + (2 * r - i + 2 * g) % g) * R + h;
          exchange(first, plane);
          rotate(lane);
          exchange(first, plane);
        }
        break;
      }
    }
  }
};

// Eigen-decomposition of one packed symmetric matrix by one warp and the density matrix of its lowest `no` eigenvectors.
//   in : S0 = the N x LD rows of the matrix (only rows / columns < m are used); Vg = this molecule's eigenvector store
//        [32][32] (row, slot) or null; warm != 0: start from the eigenvectors kept there by the previous call
//   out: S1 = rows of 2 C_occ C_occ^T (m x m valid); Vg updated; e_out[n4] / v_out[m][n4] (this molecule's, may be
//        null) in the reference's sorted order; S0 is overwritten; returns the number of sweeps
template <int N>
__device__ __forceinline__ int ring_eigh_density(double* __restrict__ S0, double* __restrict__ S1,
                                                 double* __restrict__ Vg, int warm, double tol, int m, int no, int n4,
                                                 double* __restrict__ e_out, double* __restrict__ v_out,
                                                 int check_degeneracy = 0) {
  constexpr int LD = WarpLd<N>::value;
  constexpr int R = Ring<N>::R, K = Ring<N>::K, H = Ring<N>::H;
  const int lane = threadIdx.x & 31;
  const int k = lane / R, h = lane % R;
  const int row0 = h * H;
  const int sP = 2 * k, sQ = 2 * k + 1;

  // ---- Gershgorin bounds -> shift
  double sigma;
  {
    double lo = 1.0e300, hi = -1.0e300;
    if (lane < m) {
      double rad = 0.0, aii = 0.0;
#pragma unroll
      for (int jj = 0; jj < N / 2; ++jj) {
        const double2 x = *reinterpret_cast<const double2*>(&S0[lane * LD + 2 * jj]);
        const int j0 = 2 * jj, j1 = j0 + 1;
        if (j0 < m) { if (j0 == lane) aii = x.x; else rad += fabs(x.x); }
        if (j1 < m) { if (j1 == lane) aii = x.y; else rad += fabs(x.y); }
      }
      lo = aii - rad;
      hi = aii + rad;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lo = fmin(lo, __shfl_xor_sync(FULL, lo, o));
      hi = fmax(hi, __shfl_xor_sync(FULL, hi, o));
    }
    const double w = fmax(hi - lo, 1.0);
    sigma = w - lo;      // spectrum of A + sigma I within [w, 2w]
  }

  RingCols<H, R> C;
  double (&gp)[H] = C.gp, (&gq)[H] = C.gq;
  if (warm && Vg) {
    // V0 -> S1 (row, slot), G = (A + sigma I) V0
    {
      double vv[N];
#pragma unroll
      for (int r = 0; r < N; ++r) vv[r] = lane < 2 * K ? Vg[r * WV_LD + lane] : 0.0;
#pragma unroll
      for (int r = 0; r < N; ++r)
        if (lane < 2 * K) S1[r * LD + lane] = vv[r];
    }
    __syncwarp();
#pragma unroll
    for (int i = 0; i < H; ++i) {
      const double2 x = *reinterpret_cast<const double2*>(&S1[(row0 + i) * LD + sP]);
      gp[i] = sigma * x.x;
      gq[i] = sigma * x.y;
    }
#pragma unroll 1
    for (int jj = 0; jj < N / 2; ++jj) {
      const double2 va = *reinterpret_cast<const double2*>(&S1[(2 * jj) * LD + sP]);
      const double2 vb = *reinterpret_cast<const double2*>(&S1[(2 * jj + 1) * LD + sP]);
#pragma unroll
      for (int i = 0; i < H; ++i) {
        const double2 a2 = *reinterpret_cast<const double2*>(&S0[(row0 + i) * LD + 2 * jj]);
        gp[i] = fma(a2.x, va.x, gp[i]);
        gq[i] = fma(a2.x, va.y, gq[i]);
        gp[i] = fma(a2.y, vb.x, gp[i]);
        gq[i] = fma(a2.y, vb.y, gq[i]);
      }
    }
#pragma unroll
    for (int i = 0; i < H; ++i) {      // rows beyond m carry nothing (A and V0 are zero there; keep it exact)
      if (row0 + i >= m) { gp[i] = 0.0; gq[i] = 0.0; }
    }
  } else {
#pragma unroll
    for (int i = 0; i < H; ++i) {
      const int row = row0 + i;
      const double2 x = *reinterpret_cast<const double2*>(&S0[row * LD + sP]);      // A symmetric: column = row
      gp[i] = (row < m && sP < m) ? x.x + (row == sP ? sigma : 0.0) : 0.0;
      gq[i] = (row < m && sQ < m) ? x.y + (row == sQ ? sigma : 0.0) : 0.0;
    }
  }
  __syncwarp();
  SEQM_PHASE(7);

  C.dp = C.idp = C.dq = C.idq = 1.0;
  C.ap = C.aq = 0.0;
  int sweeps = 0;
  for (; sweeps < WARP_MAX_SWEEPS;) {
    C.renorm();
    C.off = 0.0;
    const double tot = C.ap * C.ap + C.aq * C.aq;
    C.phases(Ring<N>::KA, k, h, lane);
    ++sweeps;
    // every lane of a processor carries the same off / tot
    const double off2 = 2.0 * warp_sum(C.off), tot2 = warp_sum(tot);
    if (off2 <= tol * tot2) break;
  }
  C.renorm();
  const double ap = C.ap, aq = C.aq;
  SEQM_PHASE(8);

  // ---- eigenvalues, ranking (owner lane of slot s: processor s / 2, row part s % 2)
  const bool owner = h < 2;
  const int sidx = 2 * k + (h & 1);
  const double amine = (h & 1) ? aq : ap;
  const bool real = owner && amine > 0.0;
  const double e = real ? sqrt(amine) - sigma : 0.0;
  const unsigned realmask = __ballot_sync(FULL, real);
  int rank = 0;
#pragma unroll
  for (int p = 0; p < 2 * K; ++p) {
    const int lp = (p >> 1) * R + (p & 1);
    const double ep = __shfl_sync(FULL, e, lp);
    if (((realmask >> lp) & 1u) && (ep < e || (ep == e && p < sidx))) ++rank;
  }
  SEQM_PHASE(12);
  // CHECK_DEGENERACY (diag.py:79-98, construct_P): if orbitals beyond the HOMO lie within 1e-14 of it, every orbital of
  // that degenerate set [indx1, indx2) gets the occupation (nocc - indx1) / (indx2 - indx1), the ones below it 1
  double wocc = (real && rank < no) ? 1.0 : 0.0;
  int ncw = no < m ? no : m;      // columns that carry weight
  if (check_degeneracy && no > 0 && no <= m) {
    double ehomo = 0.0;
#pragma unroll
    for (int p = 0; p < 2 * K; ++p) {
      const int lp = (p >> 1) * R + (p & 1);
      const double ep = __shfl_sync(FULL, e, lp);
      const int rp = __shfl_sync(FULL, rank, lp);
      if (((realmask >> lp) & 1u) && rp == no - 1) ehomo = ep;
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
      ncw = hi;
    }
  }
  if (e_out) {
    if (real) e_out[rank] = e;
    for (int i = m + lane; i < n4; i += 32) e_out[i] = 0.0;
  }
  const int rankP = __shfl_sync(FULL, rank, k * R), rankQ = __shfl_sync(FULL, rank, k * R + 1);
  const double wP = __shfl_sync(FULL, wocc, k * R), wQ = __shfl_sync(FULL, wocc, k * R + 1);
  const bool realP = (realmask >> (k * R)) & 1u, realQ = (realmask >> (k * R + 1)) & 1u;

  // ---- eigenvectors = normalised columns: store for the next call, reference-order output, weighted occupied
  //      columns into S0 (row, rank)
  const double nP = realP ? rsqrt(ap) : 0.0, nQ = realQ ? rsqrt(aq) : 0.0;
  for (int idx = lane; idx < N * LD; idx += 32) S0[idx] = 0.0;
  __syncwarp();
  const double fP = sqrt(wP), fQ = sqrt(wQ);
#pragma unroll
  for (int i = 0; i < H; ++i) {
    const int row = row0 + i;
    const double vp = gp[i] * nP, vq = gq[i] * nQ;
    if (Vg) *reinterpret_cast<double2*>(&Vg[row * WV_LD + sP]) = make_double2(vp, vq);
    if (v_out && row < m) {
      if (realP) v_out[(size_t)row * n4 + rankP] = vp;
      if (realQ) v_out[(size_t)row * n4 + rankQ] = vq;
    }
    if (wP > 0.0) S0[row * LD + rankP] = vp * fP;
    if (wQ > 0.0) S0[row * LD + rankQ] = vq * fQ;
  }
  __syncwarp();
  SEQM_PHASE(13);
  // ---- P = 2 W W^T, lane = row
  {
    const bool on = lane < N;
    const int rowoff = (on ? lane : 0) * LD;
    double w[N];
#pragma unroll
    for (int jj = 0; jj < N / 2; ++jj) {
      const double2 x = *reinterpret_cast<const double2*>(&S0[rowoff + 2 * jj]);
      w[2 * jj] = x.x;
      w[2 * jj + 1] = x.y;
    }
#pragma unroll 2
    for (int i2 = 0; i2 < m; ++i2) {
      double acc0 = 0.0, acc1 = 0.0;
#pragma unroll
      for (int jj = 0; jj < N / 2; ++jj) {
        if (2 * jj < ncw) {
          const double2 x = *reinterpret_cast<const double2*>(&S0[i2 * LD + 2 * jj]);
          acc0 = fma(w[2 * jj], x.x, acc0);
          acc1 = fma(w[2 * jj + 1], x.y, acc1);
        }
      }
      if (on) S1[i2 * LD + lane] = 2.0 * (acc0 + acc1);      // P[lane][i2] == P[i2][lane]
    }
  }
  __syncwarp();
  SEQM_PHASE(9);
  return sweeps;
}

}  // namespace seqm
