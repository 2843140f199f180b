// K3s: the whole SCF loop of a small molecule (packed size m <= 32) in ONE kernel launch, one warp per molecule.
// Molecules of a batch are independent (reference: every tensor of scf_loop.py carries the molecule as leading
// batch index and converged molecules are frozen by the notconverged mask), so each warp runs the reference's
// iteration schedule for its own molecule until that molecule converges: Fock build, energy, adaptive / Pulay
// mixing, Jacobi eigensolver and density all work on packed m x m matrices in shared memory; the host is not
// involved between iterations and nothing but the Pulay ring, the previous density and the eigenvector store
// (all L2 resident) goes to global memory.
// reference: seqm/seqm_functions/scf_loop.py:32-510, fock.py:6-139, energy.py:19-36, diag.py:102-172
#ifdef SEQM_SCF_TIMING
__device__ long long g_phase_acc[16];
__device__ long long g_phase_last;
#define SEQM_PHASE(i)                                                              \
  do {                                                                             \
    if (blockIdx.x == 0 && (threadIdx.x & 31) == 0) {                              \
      const long long now_ = clock64();                                            \
      g_phase_acc[(i)] += now_ - g_phase_last;                                     \
      g_phase_last = now_;                                                         \
    }                                                                              \
  } while (0)
#endif
#include <type_traits>

#include "scf_small.cuh"
#include "eigh_ring.cuh"
#include "dgemm.cuh"

namespace seqm {

__global__ void mol_atom0_kernel(int natoms, int nmol, const int32_t* __restrict__ atom_mol, int32_t* __restrict__ out) {
  const int a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a == 0) out[nmol] = natoms;
  if (a >= natoms) return;
  const int mol = atom_mol[a];
  if (a == 0 || atom_mol[a - 1] != mol) out[mol] = a;
}

template <int N, bool RING>
// Registers: up to N = 20 the shared buffers leave room for 12 and more CTAs per SM, so those sizes are compiled for 12
// (168 registers); from N = 24 on shared memory allows 7 - 10 and ptxas may take all 255 (config 2, N = 30: 0.98 ms
// against 1.03 - 1.06 ms at 168; N = 24 at 168 registers spills 556 bytes).
__global__ void __launch_bounds__(32, (N <= 20 ? 12 : 8)) scf_small_kernel(const __grid_constant__ SmallScfArgs A) {
  constexpr int LD = WarpLd<N>::value;
  __shared__ __align__(16) double Ps[N * LD];
  __shared__ __align__(16) double S0[N * LD];
  __shared__ __align__(16) double S1[N * LD];
  __shared__ __align__(16) double CS[RING ? 2 : 3 * N];
  __shared__ unsigned short TAB[RING ? 2 : N * 32];
  const seqm_topology_t& t = A.t;
  const int mol = blockIdx.x;
  const int lane = threadIdx.x;
  const int nh = A.nheavy[mol], nhy = A.nhydro[mol];
  const int nh4 = 4 * nh, m = nh4 + nhy;
  const int S = (int)t.molsize, n4 = 4 * S;
  const size_t base = (size_t)mol * n4 * n4;
  if (m <= 0) {
    if (lane == 0) { A.nc[mol] = 0; A.niter[mol] = 0; A.E[mol] = 0.0; }
    return;
  }
  const int no = A.nocc[mol];
  const int atom0 = A.mol_atom0[mol], nat = A.mol_atom0[mol + 1] - atom0;
  const int pair0 = t.start_i[atom0], pair1 = t.start_i[atom0 + nat];
  // class ranges of this molecule's pairs (the pairs of one class are consecutive in that class's slot list)
  int cls_lo[3], cls_n[3];
  {
    int lo[3] = {0x7fffffff, 0x7fffffff, 0x7fffffff}, cnt[3] = {0, 0, 0};
    for (int p = pair0 + lane; p < pair1; p += 32) {
      const int c = t.pair_cls[p], s = t.pair_slot[p];
#pragma unroll
      for (int q = 0; q < 3; ++q)
        if (c == q) { lo[q] = min(lo[q], s); ++cnt[q]; }
    }
#pragma unroll
    for (int q = 0; q < 3; ++q) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        lo[q] = min(lo[q], __shfl_xor_sync(FULL, lo[q], o));
        cnt[q] += __shfl_xor_sync(FULL, cnt[q], o);
      }
      cls_lo[q] = lo[q];
      cls_n[q] = cnt[q];
    }
  }
  double* Pold = A.Pold + (size_t)mol * A.nn;
  double* FOCK = A.FOCK ? A.FOCK + (size_t)mol * SNFOCK * A.nn : nullptr;
  double* FPPF = A.FPPF ? A.FPPF + (size_t)mol * SNFOCK * A.nn : nullptr;
  double* EMAT = A.EMAT ? A.EMAT + (size_t)mol * 36 : nullptr;
  double* Vg = A.Vstore + (size_t)mol * WV_LD * WV_LD;

#ifdef SEQM_SCF_TIMING
  if (blockIdx.x == 0 && lane == 0) g_phase_last = clock64();
#endif
  if constexpr (!RING) build_schedule<N>(TAB, lane);
  // ---- P0 -> packed
  for (int idx = lane; idx < N * LD; idx += 32) {
    const int i = idx / LD, j = idx - i * LD;
    Ps[idx] = (i < m && j < m) ? A.P[base + (size_t)pk2pd(i, nh4) * n4 + pk2pd(j, nh4)] : 0.0;
  }
  // packed, symmetrised core Hamiltonian of this molecule (hcore.py:73-158 stores the upper blocks only) -> Hpk
  double* Hpk = A.Hpk + (size_t)mol * A.nn;
  for (int idx = lane; idx < N * N; idx += 32) {
    const int i = idx / N, j = idx - i * N;
    double h = 0.0;
    if (i < m && j < m) {
      const int ai = i < nh4 ? i >> 2 : nh + i - nh4, mu = i < nh4 ? i & 3 : 0;
      const int aj = j < nh4 ? j >> 2 : nh + j - nh4, nu = j < nh4 ? j & 3 : 0;
      if (ai < aj || (ai == aj && mu <= nu)) h = A.M[(((size_t)mol * S + ai) * S + aj) * 16 + mu * 4 + nu];
      else h = A.M[(((size_t)mol * S + aj) * S + ai) * 16 + nu * 4 + mu];
    }
    Hpk[idx] = h;
  }
  for (int idx = lane; idx < N * LD; idx += 32) { S0[idx] = 0.0; S1[idx] = 0.0; }
  if (EMAT && lane == 0) {
    for (int i = 0; i < 6; ++i)
      for (int j = 0; j < 6; ++j) EMAT[i * 6 + j] = (i > j) ? -1.0 : 0.0;
  }
  __syncwarp();

  SEQM_PHASE(0);
  // F(Ps) -> S0, core Hamiltonian -> S1; returns 0.5 sum P (H + F)
  auto fock = [&]() -> double {
    SEQM_PHASE(11);
    // Descriptors of this lane's first pair in every class pass (XX, XH, XH + 32, HH): class list -> pair -> atoms ->
    // slots is a chain of three dependent L2 round trips.  Read at the head of each pass it was paid four times per
    // Fock build by a warp with nothing else to run; here the four chains go out together, level by level, and are in
    // flight under the core-Hamiltonian copy below.
    int dp[4], dsi[4], dsj[4];
    {
      int di[4], dj[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int c = q == 0 ? SEQM_CLS_XX : (q == 3 ? SEQM_CLS_HH : SEQM_CLS_XH), s = lane + (q == 2 ? 32 : 0);
        dp[q] = s < cls_n[c] ? t.cls_pairs[c][cls_lo[c] + s] : -1;
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        di[q] = dp[q] >= 0 ? t.idxi[dp[q]] : 0;
        dj[q] = dp[q] >= 0 ? t.idxj[dp[q]] : 0;
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        dsi[q] = dp[q] >= 0 ? t.atom_slot[di[q]] : 0;
        dsj[q] = dp[q] >= 0 ? t.atom_slot[dj[q]] : 0;
      }
    }
    // the same for the diagonal-block pass: atomic number, slot and the two pair ranges of this lane's atom
    int aZ = 0, aslot = 0, ai0 = 0, ai1 = 0, aj0 = 0, aj1 = 0;
    if (lane < nat) {
      const int a = atom0 + lane;
      aZ = t.Z[a]; aslot = t.atom_slot[a];
      ai0 = t.start_i[a]; ai1 = t.start_i[a + 1];
      aj0 = t.start_j[a]; aj1 = t.start_j[a + 1];
    }
    {      // all N loads of the lane in flight at once (a lone warp has nothing else to hide their latency with)
      double hv[N];
#pragma unroll
      for (int i = 0; i < N; ++i) hv[i] = lane < N ? Hpk[i * N + lane] : 0.0;
#pragma unroll
      for (int i = 0; i < N; ++i) {
        if (lane < N) {
          S1[i * LD + lane] = hv[i];
          S0[i * LD + lane] = hv[i];
        }
      }
    }
    __syncwarp();
    SEQM_PHASE(1);
    // iteration `it` of a class pass takes the prefetched descriptor a (it = 0) or b (it = 1 < nq), later ones are read here
    auto pass = [&](auto cls_tag, int nq, int pa, int sia, int sja, int pb, int sib, int sjb) {
      constexpr int CLS = decltype(cls_tag)::value;
#pragma unroll 1
      for (int s = lane, it = 0; s < cls_n[CLS]; s += 32, ++it) {
        int p, si, sj;
        if (it < nq) {
          p = it == 0 ? pa : pb;
          si = it == 0 ? sia : sib;
          sj = it == 0 ? sja : sjb;
        } else {
          p = t.cls_pairs[CLS][cls_lo[CLS] + s];
          si = t.atom_slot[t.idxi[p]];
          sj = t.atom_slot[t.idxj[p]];
        }
        small_fock_pair<CLS, LD>(A, mol, cls_lo[CLS] + s, p, si, sj, nh, nh4, Ps, S0);
      }
    };
    pass(std::integral_constant<int, SEQM_CLS_XX>{}, 1, dp[0], dsi[0], dsj[0], 0, 0, 0);
    pass(std::integral_constant<int, SEQM_CLS_XH>{}, 2, dp[1], dsi[1], dsj[1], dp[2], dsi[2], dsj[2]);
    pass(std::integral_constant<int, SEQM_CLS_HH>{}, 1, dp[3], dsi[3], dsj[3], 0, 0, 0);
    __syncwarp();
    SEQM_PHASE(2);
    // diagonal blocks: one-centre terms + the Coulomb vectors of the atom's pairs in pair order
    for (int q = lane; q < nat; q += 32) {
      const int a = atom0 + q;
      const bool pre = q == lane;      // (always, for the <= 32 atoms of a molecule this kernel takes)
      const bool heavy = (pre ? aZ : t.Z[a]) > 1;
      const int s = pre ? aslot : t.atom_slot[a];
      const int ri0 = pre ? ai0 : t.start_i[a], ri1 = pre ? ai1 : t.start_i[a + 1];
      const int rj0 = pre ? aj0 : t.start_j[a], rj1 = pre ? aj1 : t.start_j[a + 1];
      const int o = s < nh ? 4 * s : nh4 + s - nh;
      double acc[10];
#pragma unroll
      for (int k = 0; k < 10; ++k) acc[k] = 0.0;
#pragma unroll 4
      for (int r = ri0; r < ri1; ++r) {
        const double* pv = A.pairvec + (size_t)r * 20;
        acc[0] += pv[0];
        if (heavy) {
#pragma unroll
          for (int k = 1; k < 10; ++k) acc[k] += pv[k];
        }
      }
#pragma unroll 4
      for (int r = rj0; r < rj1; ++r) {
        const double* pv = A.pairvec + (size_t)t.perm_j[r] * 20 + 10;
        acc[0] += pv[0];
        if (heavy) {
#pragma unroll
          for (int k = 1; k < 10; ++k) acc[k] += pv[k];
        }
      }
      const double* Pd = Ps + o * LD + o;
      double* Fd = S0 + o * LD + o;
      const double GSS = A.gss[a], GPP = A.gpp[a], GSP = A.gsp[a], GP2 = A.gp2[a], HSP = A.hsp[a];
      if (heavy) {
        double Pu[10];
#pragma unroll
        for (int k = 0; k < 10; ++k) Pu[k] = Pd[KN(k) * LD + KM(k)];
        const double Pss = Pu[0], Ppp = Pu[2] + Pu[5] + Pu[9];
#pragma unroll
        for (int k = 0; k < 10; ++k) {
          const int mm = KM(k), n = KN(k);
          double one;
          if (k == 0) one = 0.5 * Pss * GSS + Ppp * (GSP - 0.5 * HSP);
          else if (mm == n) one = Pss * (GSP - 0.5 * HSP) + 0.5 * Pu[k] * GPP + (Ppp - Pu[k]) * (1.25 * GP2 - 0.25 * GPP);
          else if (n == 0) one = Pu[k] * (1.5 * HSP - 0.5 * GSP);
          else one = Pu[k] * (0.75 * GPP - 1.25 * GP2);
          const double f = Fd[n * LD + mm] + one + acc[k];      // Fd holds the core block
          Fd[n * LD + mm] = f;
          if (mm != n) Fd[mm * LD + n] = f;
        }
      } else {
        Fd[0] = Fd[0] + 0.5 * Pd[0] * GSS + acc[0];
      }
    }
    __syncwarp();
    SEQM_PHASE(3);
    double e = 0.0;
    for (int idx = lane; idx < m * m; idx += 32) {
      const int i = idx / m, j = idx - i * m;
      e += 0.5 * Ps[i * LD + j] * (S1[i * LD + j] + S0[i * LD + j]);
    }
    e = warp_sum(e);
    SEQM_PHASE(4);
    return e;
  };

  // P = mix(Pnew = S1, P = Ps, Pold)
  auto mix = [&](int mode) {
    double fac = 0.0;
    if (mode == SMIX_ADAPTIVE) {
      double num = 0.0, den = 0.0;
      if (lane < m) {
        const double dn = S1[lane * LD + lane], dc = Ps[lane * LD + lane], dold = Pold[lane * N + lane];
        num = (dn - dc) * (dn - dc);
        const double u = dn - dc * 2.0 + dold;
        den = u * u;
      }
      num = warp_sum(num);
      den = warp_sum(den);
      fac = sqrt(num / den);
    }
    for (int idx = lane; idx < m * m; idx += 32) {
      const int i = idx / m, j = idx - i * m;
      const double pn = S1[i * LD + j], pc = Ps[i * LD + j];
      double out;
      if (mode == SMIX_DIRECT) { Pold[i * N + j] = pc; out = pn; }
      else if (mode == SMIX_CONST) out = A.alpha * pc + (1.0 - A.alpha) * pn;
      else if (mode == SMIX_ADAPTIVE) { Pold[i * N + j] = pc; out = (1.0 + fac) * pn - fac * pc; }
      else out = pn;
      Ps[i * LD + j] = out;
    }
    __syncwarp();
  };

  // Pulay store (scf_loop.py:408-470): FOCK[c] = F, FPPF[c] = triu(FP - PF), EMAT[c][s] = <FPPF[c], FPPF[s]>
  auto pulay_store = [&](int c, int cF) {
    const bool on = lane < N;
    const int rowoff = (on ? lane : 0) * LD;
    double g[N];
#pragma unroll
    for (int j = 0; j < N; ++j) g[j] = 0.0;
#pragma unroll 2
    for (int k = 0; k < N; ++k) {
      const double fk = on ? S0[rowoff + k] : 0.0;
#pragma unroll
      for (int jj = 0; jj < N / 2; ++jj) {
        const double2 w2 = *reinterpret_cast<const double2*>(&Ps[k * LD + 2 * jj]);
        g[2 * jj] = fma(fk, w2.x, g[2 * jj]);
        g[2 * jj + 1] = fma(fk, w2.y, g[2 * jj + 1]);
      }
    }
    if (on) {
#pragma unroll
      for (int jj = 0; jj < N / 2; ++jj)
        *reinterpret_cast<double2*>(&S1[rowoff + 2 * jj]) = make_double2(g[2 * jj], g[2 * jj + 1]);
    }
    __syncwarp();
    double* fo = FOCK + (size_t)c * A.nn;
    double* fp = FPPF + (size_t)c * A.nn;
#pragma unroll
    for (int j = 0; j < N; ++j) {
      g[j] = (on && j >= lane) ? g[j] - S1[j * LD + lane] : 0.0;
      if (on) {
        fp[lane * N + j] = g[j];
        fo[lane * N + j] = S0[rowoff + j];
      }
    }
    for (int s = 0; s < cF; ++s) {
      const double* other = FPPF + (size_t)s * A.nn;
      double acc = 0.0;
      if (on && s != c) {
#pragma unroll
        for (int j = 0; j < N; ++j) acc = fma(g[j], other[lane * N + j], acc);
      } else {
#pragma unroll
        for (int j = 0; j < N; ++j) acc = fma(g[j], g[j], acc);
      }
      acc = warp_sum(acc);
      if (lane == 0) EMAT[c * 6 + s] = acc;
    }
    __syncwarp();
  };

  // F = sum_s coeff_s FOCK[s] with the DIIS coefficients of the reference (scf_loop.py:471-496) -> S0
  auto pulay_extrapolate = [&](int c, int cF) {
    // EVEC = EMAT + tril(EMAT,-1)^T, EVEC[:cF,:cF] /= EVEC[c][c], solve EVEC[:d,:d] x = e_{d-1} (d = cF + 1) by
    // Gaussian elimination with partial pivoting: lane i holds row i (7 values, static column indices), rows are
    // exchanged and broadcast with shuffles; the arithmetic and its order are those of the serial elimination
    const int d = cF + 1;
    const int row = lane < 6 ? lane : 5;
    double r[7];
#pragma unroll
    for (int j = 0; j < 6; ++j) r[j] = EMAT[row * 6 + j] + (j > row ? EMAT[j * 6 + row] : 0.0);
    {
      const double sc = EMAT[c * 6 + c];
#pragma unroll
      for (int j = 0; j < 6; ++j)
        if (row < cF && j < cF) r[j] = r[j] / sc;
    }
    r[6] = (lane == d - 1) ? 1.0 : 0.0;
#pragma unroll
    for (int cc = 0; cc < 6; ++cc) {
      if (cc < d) {
        double best = (lane >= cc && lane < d) ? fabs(r[cc]) : -1.0;
        int piv = lane;
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) {
          const double ob = __shfl_xor_sync(FULL, best, o);
          const int op = __shfl_xor_sync(FULL, piv, o);
          if (ob > best || (ob == best && op < piv)) { best = ob; piv = op; }
        }
        piv = __shfl_sync(FULL, piv, 0);
        const int src = lane == cc ? piv : (lane == piv ? cc : lane);
        double pr[7];
#pragma unroll
        for (int j = 0; j < 7; ++j) {
          r[j] = __shfl_sync(FULL, r[j], src);
          pr[j] = __shfl_sync(FULL, r[j], cc);
        }
        const double inv = 1.0 / pr[cc];
        const double f = (lane > cc && lane < d) ? r[cc] * inv : 0.0;
#pragma unroll
        for (int j = 0; j < 7; ++j)
          if (j >= cc) r[j] -= f * pr[j];
      }
    }
    double x[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
#pragma unroll
    for (int rr = 5; rr >= 0; --rr) {
      if (rr < d) {
        double sum = r[6];
#pragma unroll
        for (int j = 0; j < 6; ++j)
          if (j > rr && j < d) sum -= r[j] * x[j];
        x[rr] = __shfl_sync(FULL, sum / r[rr], rr);
      }
    }
    double coeff[SNFOCK];
#pragma unroll
    for (int s = 0; s < SNFOCK; ++s) coeff[s] = -x[s];
    constexpr int RB = (N % 6 == 0) ? 6 : 4;      // rows per batch: RB * SNFOCK loads in flight
#pragma unroll 1
    for (int i0 = 0; i0 < N; i0 += RB) {
      double fv[RB][SNFOCK];
#pragma unroll
      for (int ii = 0; ii < RB; ++ii)
#pragma unroll
        for (int s = 0; s < SNFOCK; ++s)
          fv[ii][s] = (lane < N && s < cF && i0 + ii < N) ? FOCK[(size_t)s * A.nn + (i0 + ii) * N + lane] : 0.0;
#pragma unroll
      for (int ii = 0; ii < RB; ++ii) {
        double acc = 0.0;
#pragma unroll
        for (int s = 0; s < SNFOCK; ++s)
          if (s < cF) acc += fv[ii][s] * coeff[s];
        if (lane < N && i0 + ii < N) S0[(i0 + ii) * LD + lane] = acc;
      }
    }
    __syncwarp();
  };

  // The reference's schedule (scf_loop.py: scf_forward0/1/2) as one loop with a single call site per phase.
  // k = iterations completed.  converger 2: iterations 0,1 direct, 2 adaptive, 3,4 plain (the states after
  // iterations 2 and 3 enter the Pulay ring), from 5 on the Fock matrix is the DIIS extrapolation of the ring and
  // every new state is stored.
  double E = 0.0;
  int k = -1, counter = -1, cF = 0;
  bool nc = true;
  for (;;) {
    const double en = fock();
    if (k < 0) {
      E = en;
      k = 0;
    } else {
      if (A.converger == 2 && (k == 2 || k == 3 || k >= 5)) {
        cF = cF < SNFOCK ? cF + 1 : SNFOCK;
        counter = (counter + 1) % SNFOCK;
        pulay_store(counter, cF);
        SEQM_PHASE(5);
      }
      const double err = fabs(en - E);
      E = en;
      ++k;
      nc = err > A.eps;
      if (k >= A.max_iter && (A.converger == 0 || (A.converger == 1 && k >= 3) || (A.converger == 2 && k >= 6))) break;
    }
    if (!nc) break;
    int mode;
    if (A.converger == 0) mode = SMIX_CONST;
    else if (k < 2) mode = SMIX_DIRECT;
    else if (A.converger == 1 || k == 2) mode = SMIX_ADAPTIVE;
    else mode = SMIX_TAKE;
    if (A.converger == 2 && k >= 5) pulay_extrapolate(counter, cF);
    SEQM_PHASE(6);
    if constexpr (RING) ring_eigh_density<N>(S0, S1, Vg, k > 0 ? 1 : 0, A.tol, m, no, n4, nullptr, nullptr, A.check_degeneracy);
    else warp_eigh_density<N>(S0, S1, CS, TAB, Vg, k > 0 ? 1 : 0, A.tol, m, no, n4, nullptr, nullptr, A.check_degeneracy);
    mix(mode);
    SEQM_PHASE(10);
  }

  // ---- results back to the padded layout
  for (int idx = lane; idx < m * m; idx += 32) {
    const int i = idx / m, j = idx - i * m;
    const size_t g = base + (size_t)pk2pd(i, nh4) * n4 + pk2pd(j, nh4);
    A.P[g] = Ps[i * LD + j];
    A.F[g] = S0[i * LD + j];
  }
  if (lane == 0) {
    A.E[mol] = E;
    A.nc[mol] = nc ? 1 : 0;
    A.niter[mol] = k;
    if (nc) atomicAdd(&A.stat[0], 1ULL);
    atomicMax(&A.stat[1], (unsigned long long)k);
  }
}

template <int N, bool RING>
static int launch_small_nv(int64_t nmol, const SmallScfArgs& args, cudaStream_t st) {
  static thread_local bool configured = false;
  if (!configured) {
    SEQM_CUDA(cudaFuncSetAttribute(scf_small_kernel<N, RING>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                   (int)cudaSharedmemCarveoutMaxShared));
    configured = true;
  }
  scf_small_kernel<N, RING><<<(unsigned)nmol, 32, 0, st>>>(args);
  SEQM_CHECK_LAUNCH("scf_small_kernel");
  return SEQM_OK;
}

template <int N>
static int launch_small_n(int64_t nmol, const SmallScfArgs& args, cudaStream_t st) {
  return g_eigh_warp_ring ? launch_small_nv<N, true>(nmol, args, st) : launch_small_nv<N, false>(nmol, args, st);
}


size_t scf_small_scratch_bytes(int64_t nmol) { return (size_t)(nmol + 1) * sizeof(int32_t); }

// whole SCF for a batch whose largest packed size is <= 32; stat[0..1] must be zero on entry
int launch_scf_small(const seqm_topology_t& t, seqm_w_t w, const double* M, const double* g_ss, const double* g_pp,
                     const double* g_sp, const double* g_p2, const double* h_sp, const int32_t* nheavy,
                     const int32_t* nhydro, const int32_t* nocc, int mp, int converger, double alpha, double eps,
                     int max_iter, double tol, double* P, double* F, double* E, uint8_t* nc, int32_t* niter,
                     double* pairvec, double* Pold, double* Hpk, double* FOCK, double* FPPF, double* EMAT, double* Vstore,
                     unsigned long long* stat, int32_t* mol_atom0, size_t nn, cudaStream_t st) {
  prof_mark("scf_small", st);
  mol_atom0_kernel<<<cdiv(t.natoms > 0 ? t.natoms : 1, 256), 256, 0, st>>>((int)t.natoms, (int)t.nmol, t.atom_mol, mol_atom0);
  SEQM_CHECK_LAUNCH("mol_atom0_kernel");
  SmallScfArgs a{};
  a.t = t; a.w = w; a.M = M; a.gss = g_ss; a.gpp = g_pp; a.gsp = g_sp; a.gp2 = g_p2; a.hsp = h_sp;
  a.nheavy = nheavy; a.nhydro = nhydro; a.nocc = nocc; a.mol_atom0 = mol_atom0;
  a.P = P; a.F = F; a.E = E; a.nc = nc; a.niter = niter; a.pairvec = pairvec; a.Pold = Pold; a.Hpk = Hpk; a.FOCK = FOCK;
  a.FPPF = FPPF; a.EMAT = EMAT; a.Vstore = Vstore; a.stat = stat; a.nn = nn; a.converger = converger;
  a.max_iter = max_iter; a.alpha = alpha; a.eps = eps; a.tol = tol; a.check_degeneracy = g_check_degeneracy;
  int rc;
#define SEQM_SMALL_CASE(NN) if (mp <= NN) { rc = launch_small_n<NN>(t.nmol, a, st); prof_mark("end", st); return rc; }
  SEQM_SMALL_CASE(8)
  SEQM_SMALL_CASE(12)
  SEQM_SMALL_CASE(16)
  SEQM_SMALL_CASE(20)
  SEQM_SMALL_CASE(24)
  SEQM_SMALL_CASE(28)
  SEQM_SMALL_CASE(30)
  SEQM_SMALL_CASE(32)
#undef SEQM_SMALL_CASE
  return fail(SEQM_EUNSUPPORTED, "launch_scf_small: packed size above 32");
}

}  // namespace seqm
namespace seqm {
#ifdef SEQM_SCF_TIMING
}  // namespace seqm
extern "C" int seqm_debug_phase_ticks(long long* out, int reset) {
  long long z[16] = {0};
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out, g_phase_acc, sizeof(z));
  if (reset) cudaMemcpyToSymbol(g_phase_acc, z, sizeof(z));
  return 0;
}
namespace seqm {
#endif

}  // namespace seqm
