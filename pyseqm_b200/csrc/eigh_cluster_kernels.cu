// K3a-c: sym_eig_trunc + density for packed sizes above one CTA's shared memory (118 < m <= 256): one thread-block
// CLUSTER of 8 CTAs per molecule, the matrix distributed over the cluster's shared memories (DSMEM).
//
// Method: one-sided (Hestenes) block Jacobi on G = A' V with A' = A + sigma I positive definite (sigma from the
// Gershgorin interval, so the singular values of A' are its eigenvalues and V converges to the eigenvectors of A):
//   * the mp = 32 b columns of G and V are cut into 32 column blocks of width b (b = 4, 6 or 8); CTA c of the cluster owns
//     the four block POSITIONS 4c .. 4c+3, each a column-major panel pair (G, V) in its shared memory;
//   * odd-even transposition ordering over the 32 positions: even rounds pair the positions (0,1),(2,3),.. - all local
//     to a CTA -, odd rounds pair (1,2),(3,4),..: the pair (4c+3, 4c+4) straddles two CTAs, CTA c pulls the neighbour's
//     block through distributed shared memory into a staging panel, and stores the rotated block back into the
//     neighbour's shared memory; two cluster barriers per odd round, none per even round.  A met pair swaps positions,
//     so after 32 rounds every pair of blocks has met once (one sweep);
//   * a pair: Gram matrix C = [G_I G_J]^T [G_I G_J] (2b x 2b, thread per element), its eigenvectors Q by the warp Jacobi
//     of eigh_warp.cuh (one warp per pair, two pairs per CTA in flight), its columns reordered so that Q is as close to
//     the identity as possible (needed for convergence, see jacobi_phase), [G_I G_J] <- [G_I G_J] Q and the same for V
//     (thread per matrix row, in registers);
//   * stop: the sweep in which every |c_ij| <= 1e-9 sqrt(c_ii c_jj) was the last one (quadratic convergence);
//   * eigenvalue of column k = v_k . g_k - sigma; the padding columns (unit vectors on rows >= m) never rotate and are
//     recognised by their support; rank sort (every CTA sorts all 32 b values redundantly);
//   * P = 2 C_occ C_occ^T: CTA c forms the packed rows 32c .. 32c+31: the occupied columns are pulled column by column
//     from their owners' shared memory in a fixed order (deterministic sums, no atomics).
// The inner Jacobi dominates the time (the Gram / update products are ~1/6 of a round); this kernel exists so that
// `sp2=[False]` works for C60-sized molecules like the reference (diag.py has no size limit), the fast density path at
// this size is SP2 (sp2_kernels.cu).
// reference: seqm/seqm_functions/diag.py:102-172 (sym_eig_trunc), pack.py:7-41
#include <cooperative_groups.h>

#include <mutex>

#include "dgemm.cuh"
#include "eigh_warp.cuh"

namespace cg = cooperative_groups;

namespace seqm {

constexpr int CL_NCTA = 8;            // CTAs per cluster (portable maximum)
constexpr int CL_SLOTS = 4;           // block positions per CTA
constexpr int CL_NPOS = CL_NCTA * CL_SLOTS;
constexpr int CL_NT = 256;
constexpr int CL_MAX_SWEEPS = 30;
constexpr int CL_MAX_M = 256;

struct ClusterSmem {
  // offsets in doubles from the start of dynamic shared memory
  int ld, panel, gram, cs, qmat, lam, misc, total;
};

__host__ __device__ inline ClusterSmem cluster_layout(int N) {
  const int b = N / 2, mp = CL_NPOS * b;
  ClusterSmem L;
  L.ld = mp + 1;                                    // odd column stride: conflict-free column walks
  L.panel = 0;                                      // (CL_SLOTS + 1) blocks x {G, V} x b columns x ld
  int off = (CL_SLOTS + 1) * 2 * b * L.ld;
  off = (off + 1) & ~1;
  L.gram = off; off += 2 * 2 * N * 18;              // two pairs x (S0, S1) x N rows x LD 18   (N <= 16)
  L.cs = off; off += 2 * 3 * N;
  L.qmat = off; off += 2 * N * WV_LD;
  L.lam = off; off += mp;                           // eigenvalues of all columns (every CTA holds a full copy)
  L.misc = off; off += 16 + CL_NCTA;                // sweep maximum per CTA, Gershgorin bounds, sigma
  L.total = off;
  return L;
}

// ordered-integer image of a non-negative double (for atomicMax)
__device__ __forceinline__ unsigned long long dbl_bits(double x) { return (unsigned long long)__double_as_longlong(x); }

template <int N>
__global__ void __launch_bounds__(CL_NT, 1)
eigh_cluster_kernel(int S, const double* __restrict__ F, const int32_t* __restrict__ nheavy,
                    const int32_t* __restrict__ nhydro, const int32_t* __restrict__ nocc,
                    const uint8_t* __restrict__ active, double* __restrict__ P, double* __restrict__ e_out,
                    double* __restrict__ v_out, int32_t* __restrict__ info, double tol) {
  static_assert(N == 8 || N == 12 || N == 16, "block width");
  constexpr int b = N / 2, mp = CL_NPOS * b, LDG = WarpLd<N>::value;
  static_assert(LDG == 18, "gram row stride");
  cg::cluster_group cluster = cg::this_cluster();
  const int crank = (int)cluster.block_rank();
  const int mol = blockIdx.x / CL_NCTA;
  if (active && !active[mol]) return;               // uniform over the cluster
  const int nh4 = 4 * nheavy[mol];
  const int m = nh4 + nhydro[mol];
  if (m <= 0) return;
  const int n4 = 4 * S;
  const size_t base = (size_t)mol * n4 * n4;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  extern __shared__ __align__(16) double sm[];
  const ClusterSmem L = cluster_layout(N);
  const int ld = L.ld;
  auto Gp = [&](int slot) { return sm + L.panel + (size_t)(2 * slot) * b * ld; };
  auto Vp = [&](int slot) { return sm + L.panel + (size_t)(2 * slot + 1) * b * ld; };
  double* gram = sm + L.gram;
  double* csb = sm + L.cs;
  double* qmat = sm + L.qmat;
  double* lam = sm + L.lam;
  double* misc = sm + L.misc;                       // [0..7] per-CTA values (sweep maximum / bounds), [8] sigma, [9..] scratch
  __shared__ unsigned long long s_max;
  __shared__ unsigned short TAB[N * 32];
  __shared__ int s_rank[mp];
  __shared__ unsigned char s_pad[mp];

  if (warp == 0) build_schedule<N>(TAB, lane);
  // ---- load: column j of the packed symmetric matrix (upper triangle, like eigh(UPLO='U')) into G, V = I
  double lo = 1.0e300, hi = -1.0e300;
  for (int q = 0; q < CL_SLOTS * b; ++q) {
    const int slot = q / b, col = q - slot * b;
    const int j = (CL_SLOTS * crank + slot) * b + col;          // global column index
    double* g = Gp(slot) + (size_t)col * ld;
    double* v = Vp(slot) + (size_t)col * ld;
    double rad = 0.0, ajj = 0.0;
    for (int r = tid; r < mp; r += CL_NT) {
      double a = 0.0;
      if (r < m && j < m) {
        const int rr = r <= j ? r : j, cc = r <= j ? j : r;
        a = F[base + (size_t)pk2pd(rr, nh4) * n4 + pk2pd(cc, nh4)];
      }
      g[r] = a;
      v[r] = (r == j) ? 1.0 : 0.0;
      if (r == j) ajj = a; else rad += fabs(a);
    }
    // Gershgorin interval of this column: block reduction
    __shared__ double red[2][CL_NT / 32];
    rad = warp_sum(rad);
    ajj = warp_sum(ajj);
    if (lane == 0) { red[0][warp] = rad; red[1][warp] = ajj; }
    __syncthreads();
    if (tid == 0 && j < m) {
      double R = 0.0, A = 0.0;
      for (int w = 0; w < CL_NT / 32; ++w) { R += red[0][w]; A += red[1][w]; }
      lo = fmin(lo, A - R);
      hi = fmax(hi, A + R);
    }
    __syncthreads();
  }
  // every CTA receives every CTA's interval
  if (tid == 0) {
    for (int c = 0; c < CL_NCTA; ++c) {
      double* rm = cluster.map_shared_rank(misc, c);
      rm[16 + crank] = lo;
    }
  }
  cluster.sync();
  __shared__ double s_sigma;
  double glo = 1.0e300;
  for (int c = 0; c < CL_NCTA; ++c) glo = fmin(glo, misc[16 + c]);
  cluster.sync();
  if (tid == 0) {
    for (int c = 0; c < CL_NCTA; ++c) {
      double* rm = cluster.map_shared_rank(misc, c);
      rm[16 + crank] = hi;
    }
  }
  cluster.sync();
  double ghi = -1.0e300;
  for (int c = 0; c < CL_NCTA; ++c) ghi = fmax(ghi, misc[16 + c]);
  // A' = A + sigma I with spectrum inside [spread, 2 spread]: all columns of G keep comparable norms
  const double spread = fmax(ghi - glo, 1.0);
  const double sigma = spread - glo;
  if (tid == 0) s_sigma = sigma;
  for (int q = tid; q < CL_SLOTS * b; q += CL_NT) {
    const int slot = q / b, col = q - slot * b;
    const int j = (CL_SLOTS * crank + slot) * b + col;
    Gp(slot)[(size_t)col * ld + j] += sigma;        // padding columns: sigma e_j
  }
  __syncthreads();

  // ---- one pair: Gram -> (inner Jacobi by `jwarp`) -> update.  Left block (position p) in (gl, vl), right block
  // (position p + 1) in (gr, vr); the new first half of the columns goes to the RIGHT position's storage (og1, ov1: may be
  // a neighbour's shared memory), the new second half to the LEFT position's (og2, ov2)
  auto gram_phase = [&](int pr, const double* gl, const double* gr) {
    double* C = gram + (size_t)pr * 2 * N * LDG;
    for (int e = tid; e < N * N; e += CL_NT) {
      const int i = e / N, j = e - i * N;
      const double* ci = i < b ? gl + (size_t)i * ld : gr + (size_t)(i - b) * ld;
      const double* cj = j < b ? gl + (size_t)j * ld : gr + (size_t)(j - b) * ld;
      double acc = 0.0;
#pragma unroll 4
      for (int r = 0; r < mp; ++r) acc = fma(ci[r], cj[r], acc);
      C[i * LDG + j] = acc;
    }
  };
  auto measure_phase = [&](int pr) {
    const double* C = gram + (size_t)pr * 2 * N * LDG;
    for (int e = tid; e < N * N; e += CL_NT) {
      const int i = e / N, j = e - i * N;
      if (i < j) {
        const double d = C[i * LDG + i] * C[j * LDG + j];
        const double rel = d > 0.0 ? fabs(C[i * LDG + j]) * rsqrt(d) : 0.0;
        if (rel > 0.0) atomicMax(&s_max, dbl_bits(rel));
      }
    }
  };
  auto update_phase = [&](int pr, double* gl, double* vl, double* gr, double* vr, double* og1, double* ov1, double* og2,
                          double* ov2) {
    const double* Q = gram + (size_t)pr * 2 * N * LDG;      // reordered eigenvectors of the Gram matrix (jacobi_phase)
    // 2 x mp row tasks: rows of G then rows of V
    for (int task = tid; task < 2 * mp; task += CL_NT) {
      const bool isv = task >= mp;
      const int r = isv ? task - mp : task;
      const double* inl = isv ? vl : gl;
      const double* inr = isv ? vr : gr;
      double* o1 = isv ? ov1 : og1;
      double* o2 = isv ? ov2 : og2;
      double x[N], y[N];
#pragma unroll
      for (int k = 0; k < b; ++k) { x[k] = inl[(size_t)k * ld + r]; x[b + k] = inr[(size_t)k * ld + r]; }
#pragma unroll
      for (int c = 0; c < N; ++c) y[c] = 0.0;
#pragma unroll
      for (int k = 0; k < N; ++k) {
#pragma unroll
        for (int c2 = 0; c2 < N / 2; ++c2) {
          const double2 q2 = *reinterpret_cast<const double2*>(&Q[k * LDG + 2 * c2]);
          y[2 * c2] = fma(x[k], q2.x, y[2 * c2]);
          y[2 * c2 + 1] = fma(x[k], q2.y, y[2 * c2 + 1]);
        }
      }
#pragma unroll
      for (int k = 0; k < b; ++k) { o1[(size_t)k * ld + r] = y[k]; o2[(size_t)k * ld + r] = y[b + k]; }
    }
  };
  auto jacobi_phase = [&](int pr) {      // called by one whole warp
    double* S0 = gram + (size_t)pr * 2 * N * LDG;
    double* S1 = S0 + N * LDG;
    double* Q = qmat + (size_t)pr * N * WV_LD;
    warp_eigh_density<N>(S0, S1, csb + (size_t)pr * 3 * N, TAB, Q, 0, tol, 0, 0, n4, nullptr, nullptr);
    __syncwarp();
    // The warp Jacobi leaves the eigenvectors in an arbitrary column order.  Block Jacobi only converges if the
    // transformation stays close to the identity when the pair is already nearly orthogonal (bounded-cosine condition:
    // an arbitrary column order just shuffles columns between the two blocks - tests/dev/proto/block_hestenes.py stagnates
    // at 1e-2 without this).  Greedy assignment: output column k takes the unassigned eigenvector with the largest
    // |Q[k][c]|; the reordered matrix goes to S0 (row stride LDG), where update_phase reads it.
    bool taken = lane >= N;
    int mine = 0;                               // output slot this lane's column was assigned to
    for (int k = 0; k < N; ++k) {
      double best = taken ? -1.0 : fabs(Q[k * WV_LD + lane]);
      int who = lane;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const double ob = __shfl_xor_sync(FULL, best, o);
        const int ow = __shfl_xor_sync(FULL, who, o);
        if (ob > best || (ob == best && ow < who)) { best = ob; who = ow; }
      }
      if (who == lane) { taken = true; mine = k; }
    }
    if (lane < N) {
      for (int k = 0; k < N; ++k) S0[k * LDG + mine] = Q[k * WV_LD + lane];
    }
    __syncwarp();
  };

  const int STG = CL_SLOTS;      // staging block
  int sweeps = 0;
  for (; sweeps < CL_MAX_SWEEPS; ++sweeps) {
    if (tid == 0) s_max = 0ULL;
    __syncthreads();
    for (int round = 0; round < CL_NPOS; ++round) {
      if ((round & 1) == 0) {
        gram_phase(0, Gp(0), Gp(1));
        gram_phase(1, Gp(2), Gp(3));
        __syncthreads();
        measure_phase(0);
        measure_phase(1);
        __syncthreads();
        if (warp < 2) jacobi_phase(warp);
        __syncthreads();
        update_phase(0, Gp(0), Vp(0), Gp(1), Vp(1), Gp(1), Vp(1), Gp(0), Vp(0));
        update_phase(1, Gp(2), Vp(2), Gp(3), Vp(3), Gp(3), Vp(3), Gp(2), Vp(2));
        __syncthreads();
      } else {
        const bool cross = crank + 1 < CL_NCTA;
        cluster.sync();      // the neighbour has finished writing its slot 0 (previous round)
        if (cross) {
          const double* rg = cluster.map_shared_rank(Gp(0), crank + 1);      // G and V panels of a block are adjacent
          double* stg = Gp(STG);
          for (int idx = tid; idx < 2 * b * ld; idx += CL_NT) stg[idx] = rg[idx];
        }
        __syncthreads();
        gram_phase(0, Gp(1), Gp(2));
        if (cross) gram_phase(1, Gp(3), Gp(STG));
        __syncthreads();
        measure_phase(0);
        if (cross) measure_phase(1);
        __syncthreads();
        if (warp == 0) jacobi_phase(0);
        if (warp == 1 && cross) jacobi_phase(1);
        __syncthreads();
        update_phase(0, Gp(1), Vp(1), Gp(2), Vp(2), Gp(2), Vp(2), Gp(1), Vp(1));
        if (cross) {
          double* ng = cluster.map_shared_rank(Gp(0), crank + 1);
          double* nv = cluster.map_shared_rank(Vp(0), crank + 1);
          update_phase(1, Gp(3), Vp(3), Gp(STG), Vp(STG), ng, nv, Gp(3), Vp(3));
        }
        cluster.sync();      // rotated blocks are in place on both sides
      }
    }
    // sweep maximum over the cluster
    if (tid == 0) {
      const double mine = __longlong_as_double((long long)s_max);
      for (int c = 0; c < CL_NCTA; ++c) cluster.map_shared_rank(misc, c)[crank] = mine;
    }
    cluster.sync();
    double gmax = 0.0;
    for (int c = 0; c < CL_NCTA; ++c) gmax = fmax(gmax, misc[c]);
    cluster.sync();      // everyone has read the maxima before the next sweep overwrites them
    if (gmax <= 1.0e-9) { ++sweeps; break; }
  }
  if (info && crank == 0 && tid == 0) info[mol] = sweeps;

  // ---- eigenvalues of the local columns -> every CTA
  for (int q = warp; q < CL_SLOTS * b; q += CL_NT / 32) {
    const int slot = q / b, col = q - slot * b;
    const double* g = Gp(slot) + (size_t)col * ld;
    const double* v = Vp(slot) + (size_t)col * ld;
    double dot = 0.0, padw = 0.0;
    for (int r = lane; r < mp; r += 32) {
      dot = fma(v[r], g[r], dot);
      if (r >= m) padw += fabs(v[r]);
    }
    dot = warp_sum(dot);
    padw = warp_sum(padw);
    if (lane == 0) {
      const int gid = crank * CL_SLOTS * b + q;
      const double ev = dot - s_sigma;
      // padding column: all its weight sits on rows >= m; marked by an eigenvalue image above every real one
      const double val = padw > 0.5 ? 1.0e300 : ev;
      for (int c = 0; c < CL_NCTA; ++c) cluster.map_shared_rank(lam, c)[gid] = val;
    }
  }
  cluster.sync();
  for (int i = tid; i < mp; i += CL_NT) {
    const double ei = lam[i];
    int rk = 0;
    for (int j = 0; j < mp; ++j) {
      const double ej = lam[j];
      rk += (ej < ei || (ej == ei && j < i)) ? 1 : 0;
    }
    s_rank[i] = rk;
    s_pad[i] = ei > 1.0e299 ? 1 : 0;
  }
  __syncthreads();
  const int no = min(nocc[mol], m);
  // e_out / v_out for the local columns
  for (int q = warp; q < CL_SLOTS * b; q += CL_NT / 32) {
    const int gid = crank * CL_SLOTS * b + q;
    if (s_pad[gid]) continue;
    const int rk = s_rank[gid];
    if (e_out && lane == 0) e_out[(size_t)mol * n4 + rk] = lam[gid];
    if (v_out) {
      const int slot = q / b, col = q - slot * b;
      const double* v = Vp(slot) + (size_t)col * ld;
      for (int r = lane; r < m; r += 32) v_out[base + (size_t)r * n4 + rk] = v[r];
    }
  }
  if (e_out && crank == 0)
    for (int i = m + tid; i < n4; i += CL_NT) e_out[(size_t)mol * n4 + i] = 0.0;

  // ---- density rows [32 crank, 32 crank + 32): occupied columns pulled from their owners, eight at a time
  {
    double* colbuf = Gp(STG);                 // staging block: 2 b ld doubles >= 8 ld
    constexpr int CB = 8;                     // columns per stage
    const int rloc = tid >> 3, sgrp = tid & 7;
    const int row = CL_SLOTS * b * crank + rloc;      // CL_SLOTS * b == mp / CL_NCTA rows per CTA
    constexpr int ROWS = CL_SLOTS * b;                // <= 32
    constexpr int NACC = mp / 8;
    double acc[NACC];
#pragma unroll
    for (int q = 0; q < NACC; ++q) acc[q] = 0.0;
    int nst = 0;                              // columns staged so far in the current group
    int src_cta[CB], src_off[CB];
    auto flush = [&]() {
      __syncthreads();
      for (int k = 0; k < nst; ++k) {
        const double* rv = cluster.map_shared_rank(sm + src_off[k], src_cta[k]);
        for (int r = tid; r < mp; r += CL_NT) colbuf[(size_t)k * ld + r] = rv[r];
      }
      __syncthreads();
      if (rloc < ROWS) {
        for (int k = 0; k < nst; ++k) {
          const double* cv = colbuf + (size_t)k * ld;
          const double vr = cv[row];
#pragma unroll
          for (int q = 0; q < NACC; ++q) acc[q] = fma(vr, cv[sgrp + 8 * q], acc[q]);
        }
      }
      nst = 0;
    };
    for (int gid = 0; gid < mp; ++gid) {
      if (s_pad[gid] || s_rank[gid] >= no) continue;
      const int c = gid / (CL_SLOTS * b), q = gid - c * CL_SLOTS * b;
      const int slot = q / b, col = q - slot * b;
      src_cta[nst] = c;
      src_off[nst] = L.panel + (2 * slot + 1) * b * ld + col * ld;
      if (++nst == CB) flush();
    }
    if (nst) flush();
    if (rloc < ROWS && row < m) {
      const size_t prow = base + (size_t)pk2pd(row, nh4) * n4;
#pragma unroll
      for (int q = 0; q < NACC; ++q) {
        const int s = sgrp + 8 * q;
        if (s < m) P[prow + pk2pd(s, nh4)] = 2.0 * acc[q];
      }
    }
  }
  cluster.sync();      // no CTA leaves while a peer may still read its shared memory
}

template <int N>
static int launch_cluster_n(int64_t nmol, int S, const double* F, const int32_t* nheavy, const int32_t* nhydro,
                            const int32_t* nocc, const uint8_t* active, double* P, double* e_out, double* v_out,
                            int32_t* info, double tol, cudaStream_t st) {
  const size_t smem = (size_t)cluster_layout(N).total * sizeof(double);
  static std::mutex mtx;
  static bool configured = false;
  {
    std::lock_guard<std::mutex> lock(mtx);
    if (!configured) {
      SEQM_CUDA(cudaFuncSetAttribute(eigh_cluster_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      configured = true;
    }
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(nmol * CL_NCTA), 1, 1);
  cfg.blockDim = dim3(CL_NT, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CL_NCTA;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  SEQM_CUDA(cudaLaunchKernelEx(&cfg, eigh_cluster_kernel<N>, S, F, nheavy, nhydro, nocc, active, P, e_out, v_out, info, tol));
  SEQM_CHECK_LAUNCH("eigh_cluster_kernel");
  return SEQM_OK;
}

int eigh_cluster_max_m() { return CL_MAX_M; }

int launch_eigh_cluster(int64_t nmol, int64_t molsize, int mmax, const double* F, const int32_t* nheavy,
                        const int32_t* nhydro, const int32_t* nocc, const uint8_t* active, double* P, double* e_out,
                        double* v_out, int32_t* info, double tol, cudaStream_t st) {
  if (mmax <= 4 * CL_NPOS) return launch_cluster_n<8>(nmol, (int)molsize, F, nheavy, nhydro, nocc, active, P, e_out, v_out, info, tol, st);
  if (mmax <= 6 * CL_NPOS) return launch_cluster_n<12>(nmol, (int)molsize, F, nheavy, nhydro, nocc, active, P, e_out, v_out, info, tol, st);
  if (mmax <= 8 * CL_NPOS) return launch_cluster_n<16>(nmol, (int)molsize, F, nheavy, nhydro, nocc, active, P, e_out, v_out, info, tol, st);
  return fail(SEQM_EUNSUPPORTED, "seqm_eigh_density: packed size above 256 (the cluster Jacobi limit); use the SP2 density path (sp2=[True, eps])");
}

}  // namespace seqm
