// K3a for packed sizes 32 < m <= 64: the one-sided ring Jacobi of eigh_ring.cuh on TWO warps per molecule.
// N = 4 K0 column slots on 2 K0 processors of 2 lanes (a lane holds H = N / 2 rows of its processor's columns P and Q
// in registers, 2H <= 64 doubles), K0 = 10, 12, 14 or 16 processors per warp.  A sweep of 4 K0 (K0 = 16: 63) rounds:
//   1. ring of K0 inside each warp (K0 rounds): every P of a warp meets every Q of that warp;
//   2. the warps swap their Q columns through shared memory (ONE block barrier pair), ring of K0 again: now every P
//      has met all 2 K0 Q's;
//   3. all P's go to warp 0, all Q's to warp 1 (second and last exchange through shared memory);
//   4. each warp finishes its own 2 K0 columns with the recursive phases of RingCols::phases (K0, K0 / 2, ..) without
//      any synchronisation between the warps.
// So the warps meet twice per sweep (plus once for the convergence sums); everything else is registers and shuffles.
// The shared-memory (CTA) Jacobi this replaces moved ~6 m^2 doubles through shared memory per round behind two block
// barriers: 73 % of the config-4 sweep at 13 % FP64-pipe utilisation (profiles/r01q_grad_eigh_ncu.txt).
// Shared memory: one N x (N + 2) buffer = the matrix at the start, the exchange buffer during the sweeps, the weighted
// occupied eigenvectors for P = 2 W W^T at the end.  The warm start G = (A + sigma I) V0 streams V0 from the caller's
// eigenvector store (L2) against A in shared memory.
// reference: seqm/seqm_functions/diag.py:102-172 (sym_eig_trunc, construct_P), pack.py:7-41
#include "eigh_ring.cuh"
#include "dgemm.cuh"

#ifndef SEQM_R2PHASE
#define SEQM_R2PHASE(i)      // tools/time_ring2.cu accumulates clock64() per phase of block 0
#endif

namespace seqm {

constexpr int RING2_SLOTS = 64;      // column slots = row stride of the eigenvector store of this kernel

template <int N>
__global__ void __launch_bounds__(64, (N <= 40 ? 8 : (N <= 48 ? 6 : 5)))
eigh_ring2_kernel(int S, const double* __restrict__ F, const int32_t* __restrict__ nheavy,
                  const int32_t* __restrict__ nhydro, const int32_t* __restrict__ nocc,
                  const uint8_t* __restrict__ active, double* __restrict__ P, double* __restrict__ e_out,
                  double* __restrict__ v_out, int32_t* __restrict__ info, double* __restrict__ Vstore, int warm,
                  double tol) {
  constexpr int H = N / 2, R = 2, LD = N + 2;
  constexpr int K0 = N / 4;      // processors per warp that take part (10, 12, 14, 16): 4 K0 = N column slots
  static_assert(N % 8 == 0 && N > 32 && N <= 64, "packed size handled by two warps");
  extern __shared__ __align__(16) double sm[];
  double* S0 = sm;                       // [N][LD]
  double* red = sm + N * LD;             // [8] block reductions
  double* ev = red + 8;                  // [64] eigenvalues by slot; behind it [N][8] + [N][8] staging of the warm start
  const int mol = blockIdx.x;
  if (active && !active[mol]) return;
  const int nh4 = 4 * nheavy[mol];
  const int m = nh4 + nhydro[mol];
  if (m <= 0) return;
  const int no = nocc[mol];
  const int n4 = 4 * S;
  const size_t base = (size_t)mol * n4 * n4;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int kw = lane >> 1, h = tid & 1;      // processor inside the warp, row part
  const bool idle = kw >= K0;
  const int row0 = h * H;
  // column slot of P (Q: sP + 1); the idle processors take the slots beyond the 4 K0 = N columns
  const int sP = idle ? 2 * (2 * K0 + warp * (16 - K0) + (kw - K0)) : 2 * (warp * K0 + kw);
  double* Vg = Vstore ? Vstore + (size_t)mol * N * RING2_SLOTS : nullptr;

  SEQM_R2PHASE(-1);
  // ---- the packed matrix (upper triangle, like eigh(UPLO='U')), both triangles, into shared memory
  {
    const int ctid = pk2pd(tid < m ? tid : 0, nh4);
#pragma unroll 1
    for (int r0 = 0; r0 < N; r0 += 8) {      // eight rows per batch: eight independent loads in flight per thread
      double x[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int r = r0 + u;
        x[u] = 0.0;
        if (r < m && tid < m) {
          const int cr = pk2pd(r, nh4);
          x[u] = r <= tid ? F[base + (size_t)cr * n4 + ctid] : F[base + (size_t)ctid * n4 + cr];
        }
      }
#pragma unroll
      for (int u = 0; u < 8; ++u)
        if (tid < N) S0[(r0 + u) * LD + tid] = x[u];
    }
  }
  __syncthreads();
  SEQM_R2PHASE(0);
  // ---- Gershgorin bounds -> shift
  double sigma;
  {
    double lo = 1.0e300, hi = -1.0e300;
    if (tid < m) {
      double rad = 0.0, aii = 0.0;
      for (int j = 0; j < m; ++j) {
        const double x = S0[j * LD + tid];      // column tid = row tid
        if (j == tid) aii = x; else rad += fabs(x);
      }
      lo = aii - rad;
      hi = aii + rad;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lo = fmin(lo, __shfl_xor_sync(FULL, lo, o));
      hi = fmax(hi, __shfl_xor_sync(FULL, hi, o));
    }
    if (lane == 0) { red[2 * warp] = lo; red[2 * warp + 1] = hi; }
    __syncthreads();
    lo = fmin(red[0], red[2]);
    hi = fmax(red[1], red[3]);
    const double w = fmax(hi - lo, 1.0);
    sigma = w - lo;      // spectrum of A + sigma I within [w, 2w]
  }

  SEQM_R2PHASE(1);
  // ---- G = (A + sigma I) V0, this lane's H rows of the columns P, Q of its processor
  RingCols<H, R> C;
  double (&gp)[H] = C.gp, (&gq)[H] = C.gq;
  if (warm && Vg) {
    // On the FP64 tensor pipe (DMMA m8n8k4), eight column slots at a time: both warps multiply the N x N matrix
    // (shared memory, sigma added to its diagonal) by the N x 8 chunk of V0 (staged in shared memory, the next chunk
    // prefetched into registers), the N x 8 result goes through a staging buffer to the eight lanes that own those
    // slots.  (As N FMAs per element on the FP64 pipe this product cost 38 k cycles of a ~250 k cycle solve.)
    double* vch = ev + 64;            // [N][8] chunk of V0
    double* stage = vch + N * 8;      // [N][8] chunk of G
    if (tid < m) S0[tid * LD + tid] += sigma;
#pragma unroll
    for (int i = 0; i < H; ++i) { gp[i] = 0.0; gq[i] = 0.0; }
    constexpr int NCH = N / 8, PER = (N * 8 + 63) / 64;      // chunks; V0 elements per thread and chunk
    const int fr = lane >> 2, fc = lane & 3;
    double pre[PER];
#pragma unroll
    for (int u = 0; u < PER; ++u) {
      const int e = tid + 64 * u;      // element (row e / 8, column e % 8) of the chunk
      pre[u] = e < N * 8 ? Vg[(e >> 3) * RING2_SLOTS + (e & 7)] : 0.0;
    }
#pragma unroll 1
    for (int tj = 0; tj < NCH; ++tj) {
#pragma unroll
      for (int u = 0; u < PER; ++u) {
        const int e = tid + 64 * u;
        if (e < N * 8) vch[e] = pre[u];
      }
      if (tj + 1 < NCH) {
#pragma unroll
        for (int u = 0; u < PER; ++u) {
          const int e = tid + 64 * u;
          pre[u] = e < N * 8 ? Vg[(e >> 3) * RING2_SLOTS + 8 * (tj + 1) + (e & 7)] : 0.0;
        }
      }
      __syncthreads();      // chunk of V0 (and, first time, the shifted diagonal) visible; stage free
      {
        // the row tiles warp, warp + 2, .. of this warp side by side: independent accumulators hide the DMMA latency
        constexpr int TPW = (NCH + 1) / 2;
        double c0[TPW], c1[TPW];
#pragma unroll
        for (int q = 0; q < TPW; ++q) { c0[q] = 0.0; c1[q] = 0.0; }
#pragma unroll 2
        for (int ks = 0; ks < N / 4; ++ks) {
          const double b = vch[(4 * ks + fc) * 8 + fr];
#pragma unroll
          for (int q = 0; q < TPW; ++q) {
            const int ti = warp + 2 * q;
            if (ti < NCH) {      // warp uniform
              const double a = S0[(8 * ti + fr) * LD + fc + 4 * ks];
              asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                           : "+d"(c0[q]), "+d"(c1[q])
                           : "d"(a), "d"(b));
            }
          }
        }
#pragma unroll
        for (int q = 0; q < TPW; ++q) {
          const int ti = warp + 2 * q;
          if (ti < NCH) *reinterpret_cast<double2*>(&stage[(8 * ti + fr) * 8 + 2 * fc]) = make_double2(c0[q], c1[q]);
        }
      }
      __syncthreads();
      if ((sP >> 3) == tj) {
        const int col = sP & 7;
#pragma unroll
        for (int i = 0; i < H; ++i) {
          const double2 x = *reinterpret_cast<const double2*>(&stage[(row0 + i) * 8 + col]);
          gp[i] = x.x;
          gq[i] = x.y;
        }
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
      // A symmetric: column s = row s; slots beyond m (and beyond the N rows held in shared memory) are empty
      const double xp = (row < m && sP < m) ? S0[sP * LD + row] : 0.0;
      const double xq = (row < m && sP + 1 < m) ? S0[(sP + 1) * LD + row] : 0.0;
      gp[i] = xp + ((row == sP && sP < m) ? sigma : 0.0);
      gq[i] = xq + ((row == sP + 1 && sP + 1 < m) ? sigma : 0.0);
    }
  }
  __syncthreads();      // the matrix is consumed: S0 is the exchange buffer from here on
  double* X = S0;       // [H + 3][64]

  SEQM_R2PHASE(2);
  C.dp = C.idp = C.dq = C.idq = 1.0;
  C.ap = C.aq = 0.0;
  int sweeps = 0;
  for (; sweeps < WARP_MAX_SWEEPS;) {
    C.renorm();
    C.off = 0.0;
    const double tot = C.ap * C.ap + C.aq * C.aq;
    // 1. every P of a warp meets every Q of that warp
    C.ring(K0, 0, kw, h, lane, idle);
    // 2. the warps swap their Q columns; ring again
#pragma unroll
    for (int i = 0; i < H; ++i) X[i * 64 + tid] = gq[i];
    X[H * 64 + tid] = C.dq; X[(H + 1) * 64 + tid] = C.idq; X[(H + 2) * 64 + tid] = C.aq;
    __syncthreads();
#pragma unroll
    for (int i = 0; i < H; ++i) gq[i] = X[i * 64 + (tid ^ 32)];
    C.dq = X[H * 64 + (tid ^ 32)]; C.idq = X[(H + 1) * 64 + (tid ^ 32)]; C.aq = X[(H + 2) * 64 + (tid ^ 32)];
    __syncthreads();
    C.ring(K0, 0, kw, h, lane, idle);
    // 3. all P's to warp 0 (into its Q slots), all Q's to warp 1 (into its P slots)
    if (warp == 0) {
#pragma unroll
      for (int i = 0; i < H; ++i) X[i * 64 + tid] = gq[i];
      X[H * 64 + tid] = C.dq; X[(H + 1) * 64 + tid] = C.idq; X[(H + 2) * 64 + tid] = C.aq;
    } else {
#pragma unroll
      for (int i = 0; i < H; ++i) X[i * 64 + tid] = gp[i];
      X[H * 64 + tid] = C.dp; X[(H + 1) * 64 + tid] = C.idp; X[(H + 2) * 64 + tid] = C.ap;
    }
    __syncthreads();
    if (warp == 0) {
#pragma unroll
      for (int i = 0; i < H; ++i) gq[i] = X[i * 64 + (tid ^ 32)];
      C.dq = X[H * 64 + (tid ^ 32)]; C.idq = X[(H + 1) * 64 + (tid ^ 32)]; C.aq = X[(H + 2) * 64 + (tid ^ 32)];
    } else {
#pragma unroll
      for (int i = 0; i < H; ++i) gp[i] = X[i * 64 + (tid ^ 32)];
      C.dp = X[H * 64 + (tid ^ 32)]; C.idp = X[(H + 1) * 64 + (tid ^ 32)]; C.ap = X[(H + 2) * 64 + (tid ^ 32)];
    }
    // 4. each warp finishes its own 32 columns
    C.phases(K0, kw, h, lane);
    ++sweeps;
    // every lane of a processor carries the same off / tot
    const double offw = warp_sum(C.off), totw = warp_sum(tot);
    if (lane == 0) { red[4 + 2 * warp] = offw; red[5 + 2 * warp] = totw; }
    __syncthreads();      // also: everybody has read X before the next sweep writes it
    const double off2 = 2.0 * (red[4] + red[6]), tot2 = red[5] + red[7];
    __syncthreads();
    if (off2 <= tol * tot2) break;
  }
  C.renorm();
  const double ap = C.ap, aq = C.aq;

  SEQM_R2PHASE(3);
  // ---- eigenvalues and ranking: slot tid = column P (h = 0) or Q (h = 1) of processor k
  const double amine = h ? aq : ap;
  const bool real = amine > 0.0;
  const double e = real ? sqrt(amine) - sigma : 0.0;
  ev[tid] = real ? e : 1.0e300;
  __syncthreads();
  int rank = 0;
#pragma unroll 8
  for (int p = 0; p < 64; ++p) {
    const double ep = ev[p];
    if (ep < 1.0e299 && (ep < e || (ep == e && p < tid))) ++rank;
  }
  if (e_out) {
    double* eo = e_out + (size_t)mol * n4;
    if (real) eo[rank] = e;
    for (int i = m + tid; i < n4; i += 64) eo[i] = 0.0;
  }
  const int rankP = __shfl_sync(FULL, rank, lane & ~1), rankQ = __shfl_sync(FULL, rank, lane | 1);
  const bool realP = ap > 0.0, realQ = aq > 0.0;
  const int ncw = no < m ? no : m;      // occupied columns
  SEQM_R2PHASE(4);
  // ---- eigenvectors = normalised columns: kept for the next call, reference-order output, occupied columns -> S0
  const double nP = realP ? rsqrt(ap) : 0.0, nQ = realQ ? rsqrt(aq) : 0.0;
  double* W = S0;      // [N][LD], column = rank
  for (int idx = tid; idx < N * LD; idx += 64) W[idx] = 0.0;
  __syncthreads();
  double* vo = v_out ? v_out + base : nullptr;
#pragma unroll
  for (int i = 0; i < H; ++i) {
    const int row = row0 + i;
    const double vp = gp[i] * nP, vq = gq[i] * nQ;
    if (Vg) *reinterpret_cast<double2*>(&Vg[row * RING2_SLOTS + sP]) = make_double2(vp, vq);
    if (vo && row < m) {
      if (realP) vo[(size_t)row * n4 + rankP] = vp;
      if (realQ) vo[(size_t)row * n4 + rankQ] = vq;
    }
    if (realP && rankP < ncw) W[row * LD + rankP] = vp;
    if (realQ && rankQ < ncw) W[row * LD + rankQ] = vq;
  }
  __syncthreads();
  SEQM_R2PHASE(5);
  // ---- P = 2 W W^T on the FP64 tensor pipe: 8 x 8 tiles of the upper block triangle, round robin over the two warps,
  //      DMMA m8n8k4 over the occupied columns (W is zero beyond them); a tile and its mirror image are written from the
  //      same registers, so P is exactly symmetric
  {
    constexpr int T = N / 8;
    const int fr = lane >> 2, fc = lane & 3;      // fragment row / column of this lane
    const int ksteps = (ncw + 3) >> 2;
    int tile = 0;
#pragma unroll 1
    for (int ti = 0; ti < T; ++ti) {
#pragma unroll 1
      for (int tj = ti; tj < T; ++tj, ++tile) {
        if ((tile & 1) != warp) continue;
        if (8 * ti >= m) continue;
        const double* wa = &W[(8 * ti + fr) * LD + fc];
        const double* wb = &W[(8 * tj + fr) * LD + fc];
        double c0 = 0.0, c1 = 0.0;
        for (int ks = 0; ks < ksteps; ++ks) {
          const double a = wa[4 * ks], b = wb[4 * ks];
          asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                       : "+d"(c0), "+d"(c1)
                       : "d"(a), "d"(b));
        }
        const int r = 8 * ti + fr, cA = 8 * tj + 2 * fc;
        if (r < m) {
          const size_t pr = pk2pd(r, nh4);
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            const int c = cA + u;
            if (c < m) {
              const size_t pc = pk2pd(c, nh4);
              const double val = 2.0 * (u ? c1 : c0);
              P[base + pr * n4 + pc] = val;
              if (ti != tj) P[base + pc * n4 + pr] = val;
            }
          }
        }
      }
    }
  }
  SEQM_R2PHASE(6);
  if (info && tid == 0) info[mol] = sweeps;
}

size_t eigh_ring2_vstore_doubles(int64_t nmol) { return (size_t)nmol * 64 * RING2_SLOTS; }

template <int N>
static int launch_ring2_n(int64_t nmol, int S, const double* F, const int32_t* nheavy, const int32_t* nhydro,
                          const int32_t* nocc, const uint8_t* active, double* P, double* e_out, double* v_out,
                          int32_t* info, double* Vstore, int warm, double tol, cudaStream_t st) {
  constexpr int H = N / 2, LD = N + 2;
  constexpr size_t cells = (size_t)N * LD > (size_t)64 * (H + 3) ? (size_t)N * LD : (size_t)64 * (H + 3);
  constexpr size_t smem = (cells + 8 + 64 + 2 * N * 8) * sizeof(double);
  static_assert(smem <= 48 * 1024, "fits the default dynamic shared memory limit");
  eigh_ring2_kernel<N><<<(unsigned)nmol, 64, smem, st>>>(S, F, nheavy, nhydro, nocc, active, P, e_out, v_out, info, Vstore,
                                                       warm, tol);
  SEQM_CHECK_LAUNCH("eigh_ring2_kernel");
  return SEQM_OK;
}

// mp = even padded size, 32 < mp <= 64
int launch_eigh_ring2(int64_t nmol, int64_t molsize, int mp, const double* F, const int32_t* nheavy,
                      const int32_t* nhydro, const int32_t* nocc, const uint8_t* active, double* P, double* e_out,
                      double* v_out, int32_t* info, double* Vstore, int warm, double tol, cudaStream_t st) {
#define SEQM_RING2_CASE(NN)                                                                                          \
  if (mp <= NN)                                                                                                      \
    return launch_ring2_n<NN>(nmol, (int)molsize, F, nheavy, nhydro, nocc, active, P, e_out, v_out, info, Vstore, warm, tol, st)
  SEQM_RING2_CASE(40);
  SEQM_RING2_CASE(48);
  SEQM_RING2_CASE(56);
  SEQM_RING2_CASE(64);
#undef SEQM_RING2_CASE
  return fail(SEQM_EUNSUPPORTED, "launch_eigh_ring2: packed size above 64");
}

}  // namespace seqm
