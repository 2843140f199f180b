// K3s-2w: the one-launch SCF of scf_small_kernels.cu with TWO warps (one CTA of 64 threads) per molecule.
// Same arithmetic and iteration schedule; element-wise phases stride over 64 threads, the pair pass of the Fock
// build gives every pair its own thread across both warps, matrix products and the Jacobi eigensolver split the
// columns between the warps (eigh_pair.cuh).  Config 2 has 6.9 molecules per SM: the second warp is what fills
// the issue slots the first one leaves empty.
// reference: seqm/seqm_functions/scf_loop.py:32-510, fock.py:6-139, energy.py:19-36, diag.py:102-172
#include "eigh_pair.cuh"
#include "scf_small.cuh"
#include "dgemm.cuh"

#ifndef SEQM_SMALL2_MINB
#define SEQM_SMALL2_MINB 8
#endif

namespace seqm {

template <int N>
__global__ void __launch_bounds__(64, SEQM_SMALL2_MINB) scf_small2_kernel(const __grid_constant__ SmallScfArgs A) {
  constexpr int LD = WarpLd<N>::value;
  constexpr int C = N / 2;
  __shared__ __align__(16) double Ps[N * LD];
  __shared__ __align__(16) double S0[N * LD];
  __shared__ __align__(16) double S1[N * LD];
  __shared__ __align__(16) double CS[3 * N];
  __shared__ __align__(16) double VB[64];
  __shared__ __align__(16) double red[2];
  __shared__ double coefs[8];
  __shared__ unsigned short TAB[N * 32];
  const seqm_topology_t& t = A.t;
  const int mol = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int nh = A.nheavy[mol], nhy = A.nhydro[mol];
  const int nh4 = 4 * nh, m = nh4 + nhy;
  const int S = (int)t.molsize, n4 = 4 * S;
  const size_t base = (size_t)mol * n4 * n4;
  if (m <= 0) {
    if (tid == 0) { A.nc[mol] = 0; A.niter[mol] = 0; A.E[mol] = 0.0; }
    return;
  }
  const int no = A.nocc[mol];
  const int atom0 = A.mol_atom0[mol], nat = A.mol_atom0[mol + 1] - atom0;
  const int pair0 = t.start_i[atom0], pair1 = t.start_i[atom0 + nat];
  // class ranges of this molecule's pairs (every warp computes them for itself)
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
  double* Hpk = A.Hpk + (size_t)mol * A.nn;

  if (w == 0) build_schedule<N>(TAB, lane);
  for (int idx = tid; idx < N * LD; idx += 64) {
    const int i = idx / LD, j = idx - i * LD;
    Ps[idx] = (i < m && j < m) ? A.P[base + (size_t)pk2pd(i, nh4) * n4 + pk2pd(j, nh4)] : 0.0;
    S0[idx] = 0.0;
    S1[idx] = 0.0;
  }
  // packed, symmetrised core Hamiltonian (hcore.py:73-158 stores the upper blocks only) -> Hpk
  for (int idx = tid; idx < N * N; idx += 64) {
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
  if (EMAT && tid == 0) {
    for (int i = 0; i < 6; ++i)
      for (int j = 0; j < 6; ++j) EMAT[i * 6 + j] = (i > j) ? -1.0 : 0.0;
  }
  __syncthreads();

  // F(Ps) -> S0, core Hamiltonian -> S1; returns 0.5 sum P (H + F)
  auto fock = [&]() -> double {
    for (int idx = tid; idx < N * N; idx += 64) {
      const int i = idx / N, j = idx - i * N;
      const double h = Hpk[idx];
      S1[i * LD + j] = h;
      S0[i * LD + j] = h;
    }
    __syncthreads();
    for (int s = tid; s < cls_n[2]; s += 64) small_fock_pair<SEQM_CLS_XX, LD>(A, mol, cls_lo[2] + s, nh, nh4, Ps, S0);
    for (int s = tid; s < cls_n[1]; s += 64) small_fock_pair<SEQM_CLS_XH, LD>(A, mol, cls_lo[1] + s, nh, nh4, Ps, S0);
    for (int s = tid; s < cls_n[0]; s += 64) small_fock_pair<SEQM_CLS_HH, LD>(A, mol, cls_lo[0] + s, nh, nh4, Ps, S0);
    __syncthreads();
    // diagonal blocks: thread (atom, side): side 0 sums the Coulomb vectors of the pairs in which the atom comes
    // first, side 1 those in which it comes second; the two partial sums are added in that order
    {
      const int q = tid >> 1, side = tid & 1;
      const bool has = q < nat;
      const int a = atom0 + (has ? q : 0);
      const bool heavy = t.Z[a] > 1;
      double acc[10];
#pragma unroll
      for (int k = 0; k < 10; ++k) acc[k] = 0.0;
      if (has) {
        if (side == 0) {
#pragma unroll 4
          for (int r = t.start_i[a]; r < t.start_i[a + 1]; ++r) {
            const double* pv = A.pairvec + (size_t)r * 20;
            acc[0] += pv[0];
            if (heavy) {
#pragma unroll
              for (int k = 1; k < 10; ++k) acc[k] += pv[k];
            }
          }
        } else {
#pragma unroll 4
          for (int r = t.start_j[a]; r < t.start_j[a + 1]; ++r) {
            const double* pv = A.pairvec + (size_t)t.perm_j[r] * 20 + 10;
            acc[0] += pv[0];
            if (heavy) {
#pragma unroll
              for (int k = 1; k < 10; ++k) acc[k] += pv[k];
            }
          }
        }
      }
#pragma unroll
      for (int k = 0; k < 10; ++k) {
        const double other = __shfl_xor_sync(FULL, acc[k], 1);
        acc[k] = side == 0 ? acc[k] + other : other + acc[k];
      }
      if (has && side == 0) {
        const int s = t.atom_slot[a];
        const int o = s < nh ? 4 * s : nh4 + s - nh;
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
            const double f = Fd[n * LD + mm] + one + acc[k];
            Fd[n * LD + mm] = f;
            if (mm != n) Fd[mm * LD + n] = f;
          }
        } else {
          Fd[0] = Fd[0] + 0.5 * Pd[0] * GSS + acc[0];
        }
      }
    }
    __syncthreads();
    double e = 0.0;
    for (int idx = tid; idx < m * m; idx += 64) {
      const int i = idx / m, j = idx - i * m;
      e += 0.5 * Ps[i * LD + j] * (S1[i * LD + j] + S0[i * LD + j]);
    }
    return pair_sum(e, red);
  };

  // P = mix(Pnew = S1, P = Ps, Pold)
  auto mix = [&](int mode) {
    double fac = 0.0;
    if (mode == SMIX_ADAPTIVE) {
      double num = 0.0, den = 0.0;
      if (tid < m) {
        const double dn = S1[tid * LD + tid], dc = Ps[tid * LD + tid], dold = Pold[tid * N + tid];
        num = (dn - dc) * (dn - dc);
        const double u = dn - dc * 2.0 + dold;
        den = u * u;
      }
      num = pair_sum(num, red);
      den = pair_sum(den, red);
      fac = sqrt(num / den);
    }
    for (int idx = tid; idx < m * m; idx += 64) {
      const int i = idx / m, j = idx - i * m;
      const double pn = S1[i * LD + j], pc = Ps[i * LD + j];
      double out;
      if (mode == SMIX_DIRECT) { Pold[i * N + j] = pc; out = pn; }
      else if (mode == SMIX_CONST) out = A.alpha * pc + (1.0 - A.alpha) * pn;
      else if (mode == SMIX_ADAPTIVE) { Pold[i * N + j] = pc; out = (1.0 + fac) * pn - fac * pc; }
      else out = pn;
      Ps[i * LD + j] = out;
    }
    __syncthreads();
  };

  // Pulay store (scf_loop.py:408-470): FOCK[c] = F, FPPF[c] = triu(FP - PF), EMAT[c][s] = <FPPF[c], FPPF[s]>
  auto pulay_store = [&](int c, int cF) {
    const bool on = lane < N;
    const int rowoff = (on ? lane : 0) * LD;
    const int g0 = w * C;
    double g[C];
#pragma unroll
    for (int j = 0; j < C; ++j) g[j] = 0.0;
#pragma unroll 2
    for (int k = 0; k < N; ++k) {
      const double fk = on ? S0[rowoff + k] : 0.0;
#pragma unroll
      for (int jj = 0; jj < C / 2; ++jj) {
        const double2 w2 = *reinterpret_cast<const double2*>(&Ps[k * LD + g0 + 2 * jj]);
        g[2 * jj] = fma(fk, w2.x, g[2 * jj]);
        g[2 * jj + 1] = fma(fk, w2.y, g[2 * jj + 1]);
      }
    }
    if (on) {
#pragma unroll
      for (int jj = 0; jj < C / 2; ++jj)
        *reinterpret_cast<double2*>(&S1[rowoff + g0 + 2 * jj]) = make_double2(g[2 * jj], g[2 * jj + 1]);
    }
    __syncthreads();
    double* fo = FOCK + (size_t)c * A.nn;
    double* fp = FPPF + (size_t)c * A.nn;
#pragma unroll
    for (int j = 0; j < C; ++j) {
      const int col = g0 + j;
      g[j] = (on && col >= lane) ? g[j] - S1[col * LD + lane] : 0.0;
      if (on) {
        fp[lane * N + col] = g[j];
        fo[lane * N + col] = S0[rowoff + col];
      }
    }
    for (int s = 0; s < cF; ++s) {
      const double* other = FPPF + (size_t)s * A.nn;
      double acc = 0.0;
      if (on && s != c) {
#pragma unroll
        for (int j = 0; j < C; ++j) acc = fma(g[j], other[lane * N + g0 + j], acc);
      } else {
#pragma unroll
        for (int j = 0; j < C; ++j) acc = fma(g[j], g[j], acc);
      }
      acc = pair_sum(acc, red);
      if (tid == 0) EMAT[c * 6 + s] = acc;
    }
    __syncthreads();
  };

  // F = sum_s coeff_s FOCK[s] with the DIIS coefficients of the reference (scf_loop.py:471-496) -> S0
  auto pulay_extrapolate = [&](int c, int cF) {
    if (w == 0) {
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
      if (lane == 0) {
#pragma unroll
        for (int s = 0; s < SNFOCK; ++s) coefs[s] = -x[s];
      }
    }
    __syncthreads();
    double coeff[SNFOCK];
#pragma unroll
    for (int s = 0; s < SNFOCK; ++s) coeff[s] = coefs[s];
    for (int idx = tid; idx < N * N; idx += 64) {
      const int i = idx / N, j = idx - i * N;
      double acc = 0.0;
#pragma unroll
      for (int s = 0; s < SNFOCK; ++s)
        if (s < cF) acc += FOCK[(size_t)s * A.nn + idx] * coeff[s];
      S0[i * LD + j] = acc;
    }
    __syncthreads();
  };

  // the reference's schedule as one loop with a single call site per phase (see scf_small_kernels.cu)
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
    pair_eigh_density<N>(S0, S1, CS, VB, red, TAB, Vg, k > 0 ? 1 : 0, A.tol, m, no);
    mix(mode);
  }

  // ---- results back to the padded layout
  for (int idx = tid; idx < m * m; idx += 64) {
    const int i = idx / m, j = idx - i * m;
    const size_t g = base + (size_t)pk2pd(i, nh4) * n4 + pk2pd(j, nh4);
    A.P[g] = Ps[i * LD + j];
    A.F[g] = S0[i * LD + j];
  }
  if (tid == 0) {
    A.E[mol] = E;
    A.nc[mol] = nc ? 1 : 0;
    A.niter[mol] = k;
    if (nc) atomicAdd(&A.stat[0], 1ULL);
    atomicMax(&A.stat[1], (unsigned long long)k);
  }
}

template <int N>
static int launch_small2_n(int64_t nmol, const SmallScfArgs& args, cudaStream_t st) {
  static thread_local bool configured = false;
  if (!configured) {
    SEQM_CUDA(cudaFuncSetAttribute(scf_small2_kernel<N>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                   (int)cudaSharedmemCarveoutMaxShared));
    configured = true;
  }
  scf_small2_kernel<N><<<(unsigned)nmol, 64, 0, st>>>(args);
  SEQM_CHECK_LAUNCH("scf_small2_kernel");
  return SEQM_OK;
}

// mp = even padded size (<= 32); the kernel works on the next multiple of four
int launch_scf_small2(int64_t nmol, int mp, const SmallScfArgs& a, cudaStream_t st) {
#define SEQM_SMALL2_CASE(NN) if (mp <= NN) return launch_small2_n<NN>(nmol, a, st);
  SEQM_SMALL2_CASE(8)
  SEQM_SMALL2_CASE(12)
  SEQM_SMALL2_CASE(16)
  SEQM_SMALL2_CASE(20)
  SEQM_SMALL2_CASE(24)
  SEQM_SMALL2_CASE(28)
  SEQM_SMALL2_CASE(32)
#undef SEQM_SMALL2_CASE
  return fail(SEQM_EUNSUPPORTED, "launch_scf_small2: packed size above 32");
}

}  // namespace seqm
