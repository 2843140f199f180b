// K0/K1/K4: per-atom parameter packing, the pair pass (two-centre integrals + rotation + overlap + core
// Hamiltonian + core repulsion) and its adjoint.  One thread per atom pair inside a pair CLASS
// (HH / XH / XX), so a warp never diverges on the pair type, and the integrals are stored with the pair
// index fastest ([tile][element][32]) so every global access of the warp is one 256-byte line.
//
// reference: seqm/seqm_functions/hcore.py:6-167, two_elec_two_center_int.py:7-878,
//            two_elec_two_center_int_local_frame.py:18-281, diat_overlap.py:3-393, cal_par.py:8-173,
//            energy.py:38-78; adjoint = what autograd replays for basics.py:369-388.
#include <map>
#include <mutex>
#include <vector>

#include "abi_common.cuh"
#include "pair_math.cuh"

namespace seqm {

thread_local std::string g_last_error;
void set_error(const std::string& s) { g_last_error = s; }

std::atomic<unsigned long long> g_launches{0};
static std::atomic<bool> g_prof_on{false};
static std::mutex g_prof_mutex;      // g_marks is appended to from every calling thread
struct ProfMark { std::string name; cudaEvent_t ev; };
static std::vector<ProfMark> g_marks;
void prof_mark(const char* name, cudaStream_t st) {
  if (!g_prof_on) return;
  ProfMark m;
  m.name = name;
  cudaEventCreate(&m.ev);
  cudaEventRecord(m.ev, st);
  std::lock_guard<std::mutex> lock(g_prof_mutex);
  g_marks.push_back(m);
}

unsigned long long* host_status_slot() {
  struct Slot {
    unsigned long long* p = nullptr;
    ~Slot() { if (p) cudaFreeHost(p); }
  };
  static thread_local Slot slot;
  if (!slot.p && cudaMallocHost(&slot.p, 512) != cudaSuccess) { cudaGetLastError(); slot.p = nullptr; }
  return slot.p;
}

__constant__ double c_tore[19] = {0, 1, 0, 1, 2, 3, 4, 5, 6, 7, 0, 1, 2, 3, 4, 5, 6, 7, 0};
__constant__ double c_qn[19] = {0, 1, 0, 2, 2, 2, 2, 2, 2, 2, 0, 3, 3, 3, 3, 3, 3, 3, 0};

// packed index k -> (m, n), m >= n ; IND(mu,nu) -> k
#define KM(k) ((k) == 0 ? 0 : ((k) <= 2 ? 1 : ((k) <= 5 ? 2 : 3)))
#define KN(k) ((k) - KM(k) * (KM(k) + 1) / 2)
#define IND(a, b) ((a) >= (b) ? (a) * ((a) + 1) / 2 + (b) : (b) * ((b) + 1) / 2 + (a))
#define WT(k) ((KM(k) == KN(k)) ? 1.0 : 2.0)

// ---------------------------------------------------------------------------------------------
// K0: per-atom record.  rho1/rho2 by the secant iteration of cal_par.py:51-75,138-159.
// ---------------------------------------------------------------------------------------------
template <class Fn>
__device__ double secant_solve(Fn fun, double target, double x1) {
  double x2 = x1 + 0.04;
  for (int it = 0; it < 19; ++it) {
    const double f1 = fun(x1), f2 = fun(x2);
    const double x3 = (fabs(f2 - f1) > 1.0e-16) ? x1 + (x2 - x1) * (target - f1) / (f2 - f1) : x2;
    x1 = x2;
    x2 = x3;
    if (fabs(fun(x2) - target) < 1.0e-16) break;
  }
  return x2;
}

__global__ void atom_pack_kernel(int natoms, const int32_t* __restrict__ Z, const double* __restrict__ zs,
                                 const double* __restrict__ zp, const double* __restrict__ bs,
                                 const double* __restrict__ bp, const double* __restrict__ gss,
                                 const double* __restrict__ gpp, const double* __restrict__ gp2,
                                 const double* __restrict__ hsp, const double* __restrict__ alpha,
                                 const double* __restrict__ gK, const double* __restrict__ gL,
                                 const double* __restrict__ gM, int ngauss, AtomPar* __restrict__ out) {
  const int a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= natoms) return;
  const int z = Z[a];
  AtomPar r;
  r.zs = zs[a]; r.zp = zp[a]; r.bs = bs[a]; r.bp = bp[a];
  r.alpha = alpha[a];
  r.tore = (z >= 0 && z < 19) ? c_tore[z] : 0.0;
  r.zs15 = r.zs * sqrt(r.zs);
  r.zs25 = r.zs15 * r.zs;
  r.zp25 = r.zp * r.zp * sqrt(r.zp);
  for (int k = 0; k < 4; ++k) {
    const bool on = k < ngauss;
    r.gK[k] = on ? gK[a * ngauss + k] : 0.0;
    r.gL[k] = on ? gL[a * ngauss + k] : 0.0;
    r.gM[k] = on ? gM[a * ngauss + k] : 0.0;
  }
  r.pad_[0] = r.pad_[1] = 0.0;
  r.mp.dd = r.mp.qq = r.mp.rho0 = r.mp.rho1 = r.mp.rho2 = 0.0;
  if (z == 1 || z > 2) r.mp.rho0 = 0.5 * SEQM_EV / gss[a];
  if (z > 2) {
    const double n = c_qn[z < 19 ? z : 0];
    const double D1 = (2.0 * n + 1.0) * pow(4.0 * r.zs * r.zp, n + 0.5) / pow(r.zs + r.zp, 2.0 * n + 2.0) /
                      sqrt(3.0);
    const double D2 = sqrt((4.0 * n * n + 6.0 * n + 2.0) / 20.0) / r.zp;
    r.mp.dd = D1;
    r.mp.qq = D2;
    const double h1 = hsp[a] / SEQM_EV;
    double g1 = pow(fabs(h1) / (D1 * D1), 1.0 / 3.0);
    if (h1 < 0.0) g1 = -g1;
    const double d = secant_solve(
        [D1](double x) { return 0.5 * x - 0.5 / sqrt(4.0 * D1 * D1 + 1.0 / (x * x)); }, h1, g1);
    r.mp.rho1 = 0.5 / d;
    const double h2 = 0.5 * (gpp[a] - gp2[a]) / SEQM_EV;
    double g2 = pow(fabs(h2) / 3.0 / (D2 * D2 * D2 * D2), 0.2);
    if (h2 < 0.0) g2 = -g2;
    const double q = secant_solve(
        [D2](double x) {
          return 0.25 * x - 0.5 / sqrt(4.0 * D2 * D2 + 1.0 / (x * x)) + 0.25 / sqrt(8.0 * D2 * D2 + 1.0 / (x * x));
        },
        h2, g2);
    r.mp.rho2 = 0.5 / q;
  }
  out[a] = r;
}

// ---------------------------------------------------------------------------------------------
// K1: the pair pass
// ---------------------------------------------------------------------------------------------
template <int NELEM>
struct StoreSink {
  double* w;       // class storage, already offset to this pair's lane
  double* pv;      // pairvec row of this pair
  __device__ __forceinline__ void operator()(int k, int l, double v) const {
    if (NELEM == 100) w[(size_t)(k * 10 + l) * 32] = v;
    else if (NELEM == 10) w[(size_t)k * 32] = v;
    else w[0] = v;
    if (l == 0) pv[k] = v;
    if (k == 0) pv[10 + l] = v;
  }
};

#ifndef SEQM_PAIRFWD_MINB
#define SEQM_PAIRFWD_MINB 3      // resident CTAs per SM to compile for (168 registers; measured 17 % faster than 1, tools/build_variant.sh)
#endif
constexpr int PAIRFWD_NT = 128;
// one pair (slot = its position inside the class)
template <int CLS>
__device__ __forceinline__ void pair_fwd_pair(int slot, const seqm_topology_t& t, const AtomPar* __restrict__ ap,
                                              const double* __restrict__ xij, const double* __restrict__ rij, int ngauss,
                                              double* __restrict__ wc, double* __restrict__ M,
                                              double* __restrict__ Enuc, double* __restrict__ pairvec) {
  if (slot >= t.ncls[CLS]) return;
  const int p = t.cls_pairs[CLS][slot];
  const int i = t.idxi[p], j = t.idxj[p];
  const AtomPar& a = ap[i];
  const AtomPar& b = ap[j];
  const double x[3] = {xij[3 * p], xij[3 * p + 1], xij[3 * p + 2]};
  const double r = rij[p];
  double* pv = pairvec + (size_t)p * 20;
  double gam;
  if (CLS == SEQM_CLS_XX) {
    double ri[22], R[3][3];
    repp_xx(r, a.mp, b.mp, ri);
    local_frame(x, R);
    StoreSink<100> sink{wc + ((size_t)(slot >> 5) * 100) * 32 + (slot & 31), pv};
    rotate_xx(ri, R, sink);
    gam = ri[0];
  } else if (CLS == SEQM_CLS_XH) {
    double ri[4], R[3][3];
    repp_xh(r, a.mp, b.mp, ri);
    local_frame(x, R);
    StoreSink<10> sink{wc + ((size_t)(slot >> 5) * 10) * 32 + (slot & 31), pv};
    rotate_xh(ri, R, sink);
    gam = ri[0];
  } else {
    double ri[1];
    repp_hh(r, a.mp, b.mp, ri);
    wc[slot] = ri[0];
    pv[0] = ri[0];
    pv[10] = ri[0];
    gam = ri[0];
  }
  // resonance block  M[mask] = di * (beta_i + beta_j)/2   (hcore.py:132-149)
  double di[16];
  if (r <= SEQM_OVERLAP_CUTOFF) {
    overlap_block(CLS, x, r, a, b, di);
  } else {
#pragma unroll
    for (int q = 0; q < 16; ++q) di[q] = 0.0;
  }
  const int S = (int)t.molsize;
  const size_t blk = ((size_t)t.atom_mol[i] * S + t.atom_slot[i]) * S + t.atom_slot[j];
  double* Mb = M + blk * 16;
#pragma unroll
  for (int mu = 0; mu < 4; ++mu)
#pragma unroll
    for (int nu = 0; nu < 4; ++nu)
      Mb[mu * 4 + nu] = di[mu * 4 + nu] * 0.5 * ((mu == 0 ? a.bs : a.bp) + (nu == 0 ? b.bs : b.bp));
  if (Enuc) Enuc[p] = core_repulsion(t.Z[i], t.Z[j], r, gam, a, b, ngauss);
}

template <int CLS>
__global__ void __launch_bounds__(PAIRFWD_NT, SEQM_PAIRFWD_MINB)
pair_fwd_kernel(seqm_topology_t t, const AtomPar* __restrict__ ap, const double* __restrict__ xij,
                const double* __restrict__ rij, int ngauss, double* __restrict__ wc, double* __restrict__ M,
                double* __restrict__ Enuc, double* __restrict__ pairvec) {
  pair_fwd_pair<CLS>(blockIdx.x * PAIRFWD_NT + threadIdx.x, t, ap, xij, rij, ngauss, wc, M, Enuc, pairvec);
}

// All three classes in ONE launch (CTAs [0, nbXX) heavy-heavy, the next nbXH heavy-hydrogen, the rest hydrogen-
// hydrogen) for batches whose classes each fill less than a wave: there a launch lasts as long as one thread of it, and
// three launches in a row add those three latencies up (config 2: 33 + 16 + 10 us).  Larger batches keep the separate
// launches, where the short XH / HH bodies run at their own register count and occupancy.
__global__ void __launch_bounds__(PAIRFWD_NT, SEQM_PAIRFWD_MINB)
pair_fwd_all_kernel(seqm_topology_t t, const AtomPar* __restrict__ ap, const double* __restrict__ xij,
                    const double* __restrict__ rij, int ngauss, seqm_w_t w, double* __restrict__ M,
                    double* __restrict__ Enuc, double* __restrict__ pairvec, int nbXX, int nbXH) {
  const int b = blockIdx.x;
  if (b < nbXX)
    pair_fwd_pair<SEQM_CLS_XX>(b * PAIRFWD_NT + threadIdx.x, t, ap, xij, rij, ngauss, w.wXX, M, Enuc, pairvec);
  else if (b < nbXX + nbXH)
    pair_fwd_pair<SEQM_CLS_XH>((b - nbXX) * PAIRFWD_NT + threadIdx.x, t, ap, xij, rij, ngauss, w.wXH, M, Enuc, pairvec);
  else
    pair_fwd_pair<SEQM_CLS_HH>((b - nbXX - nbXH) * PAIRFWD_NT + threadIdx.x, t, ap, xij, rij, ngauss, w.wHH, M, Enuc,
                               pairvec);
}

// diagonal blocks: U_ss/U_pp + sum over partners of the electron-core attraction, one warp per atom,
// fixed summation order (no atomics): e1b = -Z'_j w[k][0], e2a = -Z'_i w[0][k]
__global__ void __launch_bounds__(256)
hcore_diag_kernel(seqm_topology_t t, const AtomPar* __restrict__ ap, const double* __restrict__ pairvec,
                  const double* __restrict__ Uss, const double* __restrict__ Upp, double* __restrict__ M) {
  const int a = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (a >= t.natoms) return;
  const bool heavy = t.Z[a] > 1;
  double acc[10];
#pragma unroll
  for (int k = 0; k < 10; ++k) acc[k] = 0.0;
  for (int q = t.start_i[a] + lane; q < t.start_i[a + 1]; q += 32) {
    const double tj = ap[t.idxj[q]].tore;
    const double* pv = pairvec + (size_t)q * 20;
    acc[0] -= tj * pv[0];
    if (heavy) {
#pragma unroll
      for (int k = 1; k < 10; ++k) acc[k] -= tj * pv[k];
    }
  }
  for (int q = t.start_j[a] + lane; q < t.start_j[a + 1]; q += 32) {
    const int p = t.perm_j[q];
    const double ti = ap[t.idxi[p]].tore;
    const double* pv = pairvec + (size_t)p * 20 + 10;
    acc[0] -= ti * pv[0];
    if (heavy) {
#pragma unroll
      for (int k = 1; k < 10; ++k) acc[k] -= ti * pv[k];
    }
  }
#pragma unroll
  for (int k = 0; k < 10; ++k) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc[k] += __shfl_xor_sync(0xffffffffu, acc[k], o);
  }
  if (lane == 0) {
    const int S = (int)t.molsize;
    double* Mb = M + (((size_t)t.atom_mol[a] * S + t.atom_slot[a]) * S + t.atom_slot[a]) * 16;
    const double us = Uss[a], up = Upp[a];
#pragma unroll
    for (int k = 0; k < 10; ++k) {
      const int m = KM(k), n = KN(k);
      double v = acc[k];
      if (k == 0) v += us;
      else if (m == n) v += up;
      Mb[n * 4 + m] = v;
    }
  }
}

__global__ void w_expand_kernel(seqm_topology_t t, seqm_w_t w, double* __restrict__ out) {
  const long long id = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= (long long)t.npairs * 100) return;
  const int p = (int)(id / 100), e = (int)(id % 100);
  const int cls = t.pair_cls[p], slot = t.pair_slot[p];
  double v = 0.0;
  if (cls == SEQM_CLS_XX) v = w.wXX[wtile_index(slot, e, 100)];
  else if (cls == SEQM_CLS_XH) { if (e % 10 == 0) v = w.wXH[wtile_index(slot, e / 10, 10)]; }
  else { if (e == 0) v = w.wHH[slot]; }
  out[id] = v;
}

// ---------------------------------------------------------------------------------------------
// K4: adjoint of the pair pass at fixed density.  Forward-mode duals over (x1,x2,x3,r) through the very
// same templated functions (identical branches), contracted on the fly with the closed-form upstream
//   dE/dw[k][l] = qA_k pB_l + pA_k qB_l - sum Q_AB[mu][lam] P_AB[nu][sig]   (+ core attraction terms)
//   dE/dM_AB    = 2 X_AB,  with (X,Q) = (P, P/2) for the SCF energy and (D, D - P/2) for XL-BOMD.
// ---------------------------------------------------------------------------------------------
typedef Dual<4> D4;

template <int NELEM>
struct DotSink {
  const double* G;   // upstream dE/dw of this pair (local array)
  D4 acc;
  D4 w00;
  __device__ __forceinline__ void operator()(int k, int l, const D4& v) {
    const double g = (NELEM == 100) ? G[k * 10 + l] : G[k];
    acc.v += g * v.v;
#pragma unroll
    for (int q = 0; q < 4; ++q) acc.d[q] += g * v.d[q];
    if (k == 0 && l == 0) w00 = v;
  }
};

__device__ __forceinline__ void load_diag_packed(const double* __restrict__ X, size_t n4, int slot, double* out) {
  // weighted packed upper triangle of the atom's diagonal block: wt * X[n][m], k = (m,n)
  const double* base = X + ((size_t)4 * slot) * n4 + 4 * slot;
#pragma unroll
  for (int k = 0; k < 10; ++k) out[k] = WT(k) * base[(size_t)KN(k) * n4 + KM(k)];
}

#ifndef SEQM_PAIRGRAD_MINB
#define SEQM_PAIRGRAD_MINB 1
#endif
template <int CLS>
__device__ __forceinline__ void pair_grad_pair(int slot, const seqm_topology_t& t, const AtomPar* __restrict__ ap,
                                               const double* __restrict__ xij, const double* __restrict__ rij, int ngauss,
                                               const double* __restrict__ P, const double* __restrict__ Dm,
                                               const double* __restrict__ gE, const double* __restrict__ gNuc,
                                               double* __restrict__ gx, double* __restrict__ gr) {
  if (slot >= t.ncls[CLS]) return;
  const int p = t.cls_pairs[CLS][slot];
  const int i = t.idxi[p], j = t.idxj[p];
  const AtomPar& a = ap[i];
  const AtomPar& b = ap[j];
  const int S = (int)t.molsize;
  const size_t n4 = (size_t)4 * S;
  const int mol = t.atom_mol[i];
  const int si = t.atom_slot[i], sj = t.atom_slot[j];
  const double* Pm = P + (size_t)mol * n4 * n4;
  const double* Dmm = Dm ? Dm + (size_t)mol * n4 * n4 : nullptr;
  constexpr int NA = (CLS == SEQM_CLS_HH) ? 1 : 10;     // products that exist on atom i
  constexpr int NB = (CLS == SEQM_CLS_XX) ? 10 : 1;     // ... on atom j
  constexpr int OA = (CLS == SEQM_CLS_HH) ? 1 : 4;      // orbitals on atom i
  constexpr int OB = (CLS == SEQM_CLS_XX) ? 4 : 1;
  double pA[10], pB[10], qA[10], qB[10], xA[10], xB[10];
  load_diag_packed(Pm, n4, si, pA);
  load_diag_packed(Pm, n4, sj, pB);
  if (Dmm) {
    load_diag_packed(Dmm, n4, si, xA);
    load_diag_packed(Dmm, n4, sj, xB);
#pragma unroll
    for (int k = 0; k < 10; ++k) { qA[k] = xA[k] - 0.5 * pA[k]; qB[k] = xB[k] - 0.5 * pB[k]; }
  } else {
#pragma unroll
    for (int k = 0; k < 10; ++k) { xA[k] = pA[k]; xB[k] = pB[k]; qA[k] = 0.5 * pA[k]; qB[k] = 0.5 * pB[k]; }
  }
  double Pab[4][4], Qab[4][4], Xab[4][4];
#pragma unroll
  for (int mu = 0; mu < OA; ++mu)
#pragma unroll
    for (int nu = 0; nu < OB; ++nu) {
      const size_t off = ((size_t)4 * si + mu) * n4 + 4 * sj + nu;
      const double pv = Pm[off];
      Pab[mu][nu] = pv;
      if (Dmm) { const double dv = Dmm[off]; Xab[mu][nu] = dv; Qab[mu][nu] = dv - 0.5 * pv; }
      else { Xab[mu][nu] = pv; Qab[mu][nu] = 0.5 * pv; }
    }
  // upstream on the integrals
  double G[NA * NB];
#pragma unroll
  for (int k = 0; k < NA; ++k)
#pragma unroll
    for (int l = 0; l < NB; ++l) G[k * NB + l] = qA[k] * pB[l] + pA[k] * qB[l];
#pragma unroll
  for (int mu = 0; mu < OA; ++mu)
#pragma unroll
    for (int nu = 0; nu < OA; ++nu)
#pragma unroll
      for (int la = 0; la < OB; ++la)
#pragma unroll
        for (int sg = 0; sg < OB; ++sg) G[IND(mu, nu) * NB + IND(la, sg)] -= Qab[mu][la] * Pab[nu][sg];
#pragma unroll
  for (int k = 0; k < NA; ++k) G[k * NB] -= b.tore * xA[k];
#pragma unroll
  for (int l = 0; l < NB; ++l) G[l] -= a.tore * xB[l];

  D4 x[3], r;
#pragma unroll
  for (int c = 0; c < 3; ++c) { x[c] = D4(xij[3 * p + c]); x[c].d[c] = 1.0; }
  r = D4(rij[p]);
  r.d[3] = 1.0;

  D4 eel(0.0), gam;
  if (CLS == SEQM_CLS_XX) {
    D4 ri[22], R[3][3];
    repp_xx(r, a.mp, b.mp, ri);
    local_frame(x, R);
    DotSink<100> sink{G, D4(0.0), D4(0.0)};
    rotate_xx(ri, R, sink);
    eel = sink.acc; gam = sink.w00;
  } else if (CLS == SEQM_CLS_XH) {
    D4 ri[4], R[3][3];
    repp_xh(r, a.mp, b.mp, ri);
    local_frame(x, R);
    DotSink<10> sink{G, D4(0.0), D4(0.0)};
    rotate_xh(ri, R, sink);
    eel = sink.acc; gam = sink.w00;
  } else {
    D4 ri[1];
    repp_hh(r, a.mp, b.mp, ri);
    eel = ri[0] * G[0];
    gam = ri[0];
  }
  if (rij[p] <= SEQM_OVERLAP_CUTOFF) {
    D4 di[16];
    overlap_block(CLS, x, r, a, b, di);
#pragma unroll
    for (int mu = 0; mu < OA; ++mu)
#pragma unroll
      for (int nu = 0; nu < OB; ++nu)
        eel = eel + di[mu * 4 + nu] * (Xab[mu][nu] * ((mu == 0 ? a.bs : a.bp) + (nu == 0 ? b.bs : b.bp)));
  }
  const double ge = gE ? gE[mol] : 1.0;
  const double gn = gNuc ? gNuc[p] : 1.0;
  const D4 en = core_repulsion(t.Z[i], t.Z[j], r, gam, a, b, ngauss);
  gx[3 * p + 0] = ge * eel.d[0] + gn * en.d[0];
  gx[3 * p + 1] = ge * eel.d[1] + gn * en.d[1];
  gx[3 * p + 2] = ge * eel.d[2] + gn * en.d[2];
  gr[p] = ge * eel.d[3] + gn * en.d[3];
}

template <int CLS>
__global__ void __launch_bounds__(128, SEQM_PAIRGRAD_MINB)
pair_grad_kernel(seqm_topology_t t, const AtomPar* __restrict__ ap, const double* __restrict__ xij,
                 const double* __restrict__ rij, int ngauss, const double* __restrict__ P,
                 const double* __restrict__ Dm, const double* __restrict__ gE, const double* __restrict__ gNuc,
                 double* __restrict__ gx, double* __restrict__ gr) {
  pair_grad_pair<CLS>(blockIdx.x * blockDim.x + threadIdx.x, t, ap, xij, rij, ngauss, P, Dm, gE, gNuc, gx, gr);
}



// the classes of a pair pass share one launch while all their CTAs together stay within this many waves of resident CTAs
constexpr int PAIR_MERGE_WAVES = 4;
static int g_pair_grad_reverse = 1;      // 0: forward-mode kernel for heavy-heavy pairs too (test / comparison hook)

// ---------------------------------------------------------------------------------------------
// K4 for heavy-heavy pairs, reverse mode (replaces pair_grad_kernel<XX>, which pushed four tangents through the
// integrals AND the rotation with the 100-element upstream in registers: 7.9 KB of spills per thread).
//   * the upstream G[10][10] of the pair lives in shared memory, element-major (conflict-free, one column per thread)
//   * local integrals with ONE tangent (d/dr, Dual<1>)
//   * rotate_xx_adjoint (pair_math.cuh): one pass over the rows of G in plain doubles gives dE/dr and the adjoint of
//     the rotation matrix; the three tangents of the local frame turn that into dE/dx
//   * overlap (four tangents, small) and core repulsion (one tangent) as before
// ---------------------------------------------------------------------------------------------
constexpr int PGX_NT = 32;

struct SmemG {
  const double* g;      // this thread's column: element e at g[e * PGX_NT]
  __device__ __forceinline__ double operator()(int k, int l) const { return g[(k * 10 + l) * PGX_NT]; }
};

#ifndef SEQM_PGX_MINB
#define SEQM_PGX_MINB 1      // 255 registers, two warps per scheduler (three would need <= 168 registers: 2.7 KB of spills)
#endif
// one heavy-heavy pair; Gs: the CTA's [100][PGX_NT] upstream buffer in shared memory
__device__ __forceinline__ void pair_grad_xx_pair(int slot, double* Gs, const seqm_topology_t& t,
                                                  const AtomPar* __restrict__ ap, const double* __restrict__ xij,
                                                  const double* __restrict__ rij, int ngauss, const double* __restrict__ P,
                                                  const double* __restrict__ Dm, const double* __restrict__ gE,
                                                  const double* __restrict__ gNuc, double* __restrict__ gx,
                                                  double* __restrict__ gr) {
  if (slot >= t.ncls[SEQM_CLS_XX]) return;
  const int p = t.cls_pairs[SEQM_CLS_XX][slot];
  const int i = t.idxi[p], j = t.idxj[p];
  const AtomPar& a = ap[i];
  const AtomPar& b = ap[j];
  const int S = (int)t.molsize;
  const size_t n4 = (size_t)4 * S;
  const int mol = t.atom_mol[i];
  const int si = t.atom_slot[i], sj = t.atom_slot[j];
  const double* Pm = P + (size_t)mol * n4 * n4;
  const double* Dmm = Dm ? Dm + (size_t)mol * n4 * n4 : nullptr;
  double* Gt = Gs + threadIdx.x;
  double eov[4] = {0.0, 0.0, 0.0, 0.0};      // overlap part of dE/d(x1, x2, x3, r)
  {
    double pA[10], pB[10], qA[10], qB[10], xA[10], xB[10];
    load_diag_packed(Pm, n4, si, pA);
    load_diag_packed(Pm, n4, sj, pB);
    if (Dmm) {
      load_diag_packed(Dmm, n4, si, xA);
      load_diag_packed(Dmm, n4, sj, xB);
#pragma unroll
      for (int k = 0; k < 10; ++k) { qA[k] = xA[k] - 0.5 * pA[k]; qB[k] = xB[k] - 0.5 * pB[k]; }
    } else {
#pragma unroll
      for (int k = 0; k < 10; ++k) { xA[k] = pA[k]; xB[k] = pB[k]; qA[k] = 0.5 * pA[k]; qB[k] = 0.5 * pB[k]; }
    }
    double Pab[4][4], Qab[4][4];
#pragma unroll
    for (int mu = 0; mu < 4; ++mu)
#pragma unroll
      for (int nu = 0; nu < 4; ++nu) {
        const size_t off = ((size_t)4 * si + mu) * n4 + 4 * sj + nu;
        const double pv = Pm[off];
        Pab[mu][nu] = pv;
        Qab[mu][nu] = Dmm ? Dmm[off] - 0.5 * pv : 0.5 * pv;
      }
    // upstream on the integrals, one element at a time: Coulomb, exchange (every ordering of the two index pairs),
    // and the core-electron attraction terms that ride on the (mu nu|ss) / (ss|la sg) integrals
#pragma unroll
    for (int mu = 0; mu < 4; ++mu)
#pragma unroll
      for (int nu = 0; nu <= mu; ++nu) {
        const int k = IND(mu, nu);
#pragma unroll
        for (int la = 0; la < 4; ++la)
#pragma unroll
          for (int sg = 0; sg <= la; ++sg) {
            const int l = IND(la, sg);
            double g = qA[k] * pB[l] + pA[k] * qB[l] - Qab[mu][la] * Pab[nu][sg];
            if (la != sg) g -= Qab[mu][sg] * Pab[nu][la];
            if (mu != nu) {
              g -= Qab[nu][la] * Pab[mu][sg];
              if (la != sg) g -= Qab[nu][sg] * Pab[mu][la];
            }
            if (l == 0) g -= b.tore * xA[k];
            if (k == 0) g -= a.tore * xB[l];
            Gt[(k * 10 + l) * PGX_NT] = g;
          }
      }
  }
  // Compiler fence: without it nvcc forwards the 100 stores above to the reads of rotate_xx_adjoint (same thread,
  // static indices), i.e. keeps the whole upstream in registers after all: the SASS had 100 STS, not one LDS, and
  // 2.2 KB of spills (0.58 ms for 256 C60; 0.38 ms with the fence and the overlap adjoint below).
  asm volatile("" ::: "memory");
  {
    // overlap part: dE/dM block = 2 X_AB (X = P, or D in the XL-BOMD energy), M = overlap * mean beta; reverse mode
    // (overlap_xx_adjoint, pair_math.cuh: radial integrals with one tangent, angular factors with three)
    if (rij[p] <= SEQM_OVERLAP_CUTOFF) {
      const double* Xm = Dmm ? Dmm : Pm;
      double F[16];
#pragma unroll
      for (int mu = 0; mu < 4; ++mu)
#pragma unroll
        for (int nu = 0; nu < 4; ++nu)
          F[mu * 4 + nu] = Xm[((size_t)4 * si + mu) * n4 + 4 * sj + nu] * ((mu == 0 ? a.bs : a.bp) + (nu == 0 ? b.bs : b.bp));
      const double xo[3] = {xij[3 * p], xij[3 * p + 1], xij[3 * p + 2]};
      overlap_xx_adjoint(xo, rij[p], a, b, F, eov);
    }
  }
  typedef Dual<1> D1;
  D1 r1(rij[p]);
  r1.d[0] = 1.0;
  double ri[22], dri[22];
  {
    D1 ri1[22];
    repp_xx(r1, a.mp, b.mp, ri1);
#pragma unroll
    for (int q = 0; q < 22; ++q) { ri[q] = ri1[q].v; dri[q] = ri1[q].d[0]; }
  }
  const double xv[3] = {xij[3 * p], xij[3 * p + 1], xij[3 * p + 2]};
  double grot[3], grr;
  rotate_xx_adjoint(ri, dri, xv, SmemG{Gt}, grot, grr);
  D1 gam(ri[0]);
  gam.d[0] = dri[0];
  const D1 en = core_repulsion(t.Z[i], t.Z[j], r1, gam, a, b, ngauss);
  const double ge = gE ? gE[mol] : 1.0;
  const double gn = gNuc ? gNuc[p] : 1.0;
  gx[3 * p + 0] = ge * (grot[0] + eov[0]);
  gx[3 * p + 1] = ge * (grot[1] + eov[1]);
  gx[3 * p + 2] = ge * (grot[2] + eov[2]);
  gr[p] = ge * (grr + eov[3]) + gn * en.d[0];
}

__global__ void __launch_bounds__(PGX_NT, SEQM_PGX_MINB)
pair_grad_xx_kernel(seqm_topology_t t, const AtomPar* __restrict__ ap, const double* __restrict__ xij,
                    const double* __restrict__ rij, int ngauss, const double* __restrict__ P,
                    const double* __restrict__ Dm, const double* __restrict__ gE, const double* __restrict__ gNuc,
                    double* __restrict__ gx, double* __restrict__ gr) {
  __shared__ double Gs[100 * PGX_NT];
  pair_grad_xx_pair(blockIdx.x * PGX_NT + threadIdx.x, Gs, t, ap, xij, rij, ngauss, P, Dm, gE, gNuc, gx, gr);
}

// All three classes in one launch of one-warp CTAs (see pair_fwd_all_kernel): [0, nbXX) heavy-heavy in reverse mode, then
// heavy-hydrogen and hydrogen-hydrogen with forward duals.
__global__ void __launch_bounds__(PGX_NT, SEQM_PGX_MINB)
pair_grad_all_kernel(seqm_topology_t t, const AtomPar* __restrict__ ap, const double* __restrict__ xij,
                     const double* __restrict__ rij, int ngauss, const double* __restrict__ P,
                     const double* __restrict__ Dm, const double* __restrict__ gE, const double* __restrict__ gNuc,
                     double* __restrict__ gx, double* __restrict__ gr, int nbXX, int nbXH) {
  __shared__ double Gs[100 * PGX_NT];
  const int b = blockIdx.x;
  if (b < nbXX)
    pair_grad_xx_pair(b * PGX_NT + threadIdx.x, Gs, t, ap, xij, rij, ngauss, P, Dm, gE, gNuc, gx, gr);
  else if (b < nbXX + nbXH)
    pair_grad_pair<SEQM_CLS_XH>((b - nbXX) * PGX_NT + threadIdx.x, t, ap, xij, rij, ngauss, P, Dm, gE, gNuc, gx, gr);
  else
    pair_grad_pair<SEQM_CLS_HH>((b - nbXX - nbXH) * PGX_NT + threadIdx.x, t, ap, xij, rij, ngauss, P, Dm, gE, gNuc, gx, gr);
}

// ---------------------------------------------------------------------------------------------
// Pair geometry on the device (the arithmetic of the reference Parser, basics.py:91-108) and its adjoint:
//   d = R_j - R_i, dist = sqrt(d.d) (1e-4 A^2 if zero), xij = d/dist, rij = dist * lcf
//   grad R[atom] = sum over its pairs of +-( (g_x - (g_x . x) x)/dist + g_r lcf x ), fixed order, no atomics
// ---------------------------------------------------------------------------------------------
__global__ void pair_geometry_kernel(seqm_topology_t t, const double* __restrict__ R, double lcf,
                                     double* __restrict__ xij, double* __restrict__ rij) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= t.npairs) return;
  const int i = t.idxi[p], j = t.idxj[p];
  const int S = (int)t.molsize;
  const double* ri = R + ((size_t)t.atom_mol[i] * S + t.atom_slot[i]) * 3;
  const double* rj = R + ((size_t)t.atom_mol[j] * S + t.atom_slot[j]) * 3;
  // the reference's arithmetic operation by operation, no FMA contraction: bit-identical to basics.py:91-108
  const double dx = __dsub_rn(rj[0], ri[0]), dy = __dsub_rn(rj[1], ri[1]), dz = __dsub_rn(rj[2], ri[2]);
  double d2 = __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
  if (d2 == 0.0) d2 = 1.0e-4;
  const double dist = __dsqrt_rn(d2);
  xij[3 * p] = __ddiv_rn(dx, dist); xij[3 * p + 1] = __ddiv_rn(dy, dist); xij[3 * p + 2] = __ddiv_rn(dz, dist);
  rij[p] = __dmul_rn(dist, lcf);
}

__device__ __forceinline__ void pair_geom_grad_one(const double* __restrict__ xij, const double* __restrict__ rij,
                                                   const double* __restrict__ gx, const double* __restrict__ gr,
                                                   double lcf, int p, double* g) {
  const double x0 = xij[3 * p], x1 = xij[3 * p + 1], x2 = xij[3 * p + 2];
  const double g0 = gx[3 * p], g1 = gx[3 * p + 1], g2 = gx[3 * p + 2];
  const double inv = lcf / rij[p];                  // 1/dist
  const double dot = g0 * x0 + g1 * x1 + g2 * x2;
  const double gd = gr[p] * lcf;
  g[0] = (g0 - dot * x0) * inv + gd * x0;
  g[1] = (g1 - dot * x1) * inv + gd * x1;
  g[2] = (g2 - dot * x2) * inv + gd * x2;
}

__global__ void __launch_bounds__(256)
pair_geometry_grad_kernel(seqm_topology_t t, const double* __restrict__ xij, const double* __restrict__ rij,
                          const double* __restrict__ gx, const double* __restrict__ gr, double lcf,
                          double* __restrict__ gR) {
  const int a = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (a >= t.natoms) return;
  double acc[3] = {0.0, 0.0, 0.0}, g[3];
  for (int q = t.start_i[a] + lane; q < t.start_i[a + 1]; q += 32) {      // atom is i: d = R_j - R_i
    pair_geom_grad_one(xij, rij, gx, gr, lcf, q, g);
    acc[0] -= g[0]; acc[1] -= g[1]; acc[2] -= g[2];
  }
  for (int q = t.start_j[a] + lane; q < t.start_j[a + 1]; q += 32) {
    pair_geom_grad_one(xij, rij, gx, gr, lcf, t.perm_j[q], g);
    acc[0] += g[0]; acc[1] += g[1]; acc[2] += g[2];
  }
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc[c] += __shfl_xor_sync(0xffffffffu, acc[c], o);
  if (lane == 0) {
    double* out = gR + ((size_t)t.atom_mol[a] * t.molsize + t.atom_slot[a]) * 3;
    out[0] = acc[0]; out[1] = acc[1]; out[2] = acc[2];
  }
}

}  // namespace seqm

using namespace seqm;

extern "C" {

const char* seqm_last_error(void) { return g_last_error.c_str(); }

unsigned long long seqm_launch_count(void) { return g_launches.load(); }

void seqm_profile_enable(int on) {
  std::lock_guard<std::mutex> lock(g_prof_mutex);
  for (auto& m : g_marks) cudaEventDestroy(m.ev);
  g_marks.clear();
  g_prof_on = on != 0;
}

// "name total_ms count\n" per region name; synchronises the device
int seqm_profile_report(char* buf, size_t cap) {
  cudaDeviceSynchronize();
  std::lock_guard<std::mutex> lock(g_prof_mutex);
  std::map<std::string, std::pair<double, int>> acc;
  for (size_t i = 0; i + 1 < g_marks.size(); ++i) {
    if (g_marks[i].name == "end") continue;
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, g_marks[i].ev, g_marks[i + 1].ev) != cudaSuccess) { cudaGetLastError(); continue; }
    auto& a = acc[g_marks[i].name];
    a.first += ms;
    a.second += 1;
  }
  std::string out;
  for (auto& kv : acc) {
    char line[256];
    snprintf(line, sizeof line, "%s %.6f %d\n", kv.first.c_str(), kv.second.first, kv.second.second);
    out += line;
  }
  if (buf && cap) { snprintf(buf, cap, "%s", out.c_str()); }
  return (int)out.size();
}
int seqm_version(void) { return 100; }
size_t seqm_atompar_bytes(void) { return sizeof(AtomPar); }

int seqm_atom_pack(int64_t natoms, const int32_t* Z, const double* zeta_s, const double* zeta_p,
                   const double* beta_s, const double* beta_p, const double* g_ss, const double* g_pp,
                   const double* g_p2, const double* h_sp, const double* alpha, const double* gK,
                   const double* gL, const double* gM, int ngauss, void* atompar_out, void* stream) {
  SEQM_REQUIRE(natoms >= 0 && ngauss >= 0 && ngauss <= 4, "seqm_atom_pack: bad sizes");
  SEQM_REQUIRE(Z && zeta_s && zeta_p && beta_s && beta_p && g_ss && g_pp && g_p2 && h_sp && alpha && atompar_out,
               "seqm_atom_pack: null pointer");
  SEQM_REQUIRE(ngauss == 0 || (gK && gL && gM), "seqm_atom_pack: gaussians missing");
  if (natoms == 0) return SEQM_OK;
  atom_pack_kernel<<<cdiv(natoms, 128), 128, 0, (cudaStream_t)stream>>>(
      (int)natoms, Z, zeta_s, zeta_p, beta_s, beta_p, g_ss, g_pp, g_p2, h_sp, alpha, gK, gL, gM, ngauss,
      (AtomPar*)atompar_out);
  SEQM_CHECK_LAUNCH("atom_pack_kernel");
  return SEQM_OK;
}

int seqm_hcore(const seqm_topology_t* topo, const void* atompar, const double* xij, const double* rij,
               const double* U_ss, const double* U_pp, int ngauss, seqm_w_t w, double* M, double* EnucAB,
               double* pairvec, void* stream) {
  SEQM_REQUIRE(topo && atompar && xij && rij && U_ss && U_pp && M && pairvec, "seqm_hcore: null pointer");
  SEQM_REQUIRE(ngauss >= 0 && ngauss <= 4, "seqm_hcore: ngauss");
  const seqm_topology_t& t = *topo;
  cudaStream_t st = (cudaStream_t)stream;
  const AtomPar* ap = (const AtomPar*)atompar;
  prof_mark("pair_fwd", st);
  SEQM_REQUIRE(t.ncls[SEQM_CLS_XX] == 0 || w.wXX, "seqm_hcore: wXX null");
  SEQM_REQUIRE(t.ncls[SEQM_CLS_XH] == 0 || w.wXH, "seqm_hcore: wXH null");
  SEQM_REQUIRE(t.ncls[SEQM_CLS_HH] == 0 || w.wHH, "seqm_hcore: wHH null");
  const long long nbXX = cdiv(t.ncls[2], PAIRFWD_NT), nbXH = cdiv(t.ncls[1], PAIRFWD_NT), nbHH = cdiv(t.ncls[0], PAIRFWD_NT);
  const int nclasses = (nbXX > 0) + (nbXH > 0) + (nbHH > 0);
  if (nclasses > 1 && nbXX + nbXH + nbHH <= PAIR_MERGE_WAVES * 148 * SEQM_PAIRFWD_MINB) {
    pair_fwd_all_kernel<<<(unsigned)(nbXX + nbXH + nbHH), PAIRFWD_NT, 0, st>>>(t, ap, xij, rij, ngauss, w, M, EnucAB, pairvec,
                                                                             (int)nbXX, (int)nbXH);
    SEQM_CHECK_LAUNCH("pair_fwd_all_kernel");
  } else {
    if (nbXX > 0) {
      pair_fwd_kernel<SEQM_CLS_XX><<<(unsigned)nbXX, PAIRFWD_NT, 0, st>>>(t, ap, xij, rij, ngauss, w.wXX, M, EnucAB, pairvec);
      SEQM_CHECK_LAUNCH("pair_fwd_kernel<XX>");
    }
    if (nbXH > 0) {
      pair_fwd_kernel<SEQM_CLS_XH><<<(unsigned)nbXH, PAIRFWD_NT, 0, st>>>(t, ap, xij, rij, ngauss, w.wXH, M, EnucAB, pairvec);
      SEQM_CHECK_LAUNCH("pair_fwd_kernel<XH>");
    }
    if (nbHH > 0) {
      pair_fwd_kernel<SEQM_CLS_HH><<<(unsigned)nbHH, PAIRFWD_NT, 0, st>>>(t, ap, xij, rij, ngauss, w.wHH, M, EnucAB, pairvec);
      SEQM_CHECK_LAUNCH("pair_fwd_kernel<HH>");
    }
  }
  prof_mark("hcore_diag", st);
  if (t.natoms > 0) {
    hcore_diag_kernel<<<cdiv(t.natoms * 32, 256), 256, 0, st>>>(t, ap, pairvec, U_ss, U_pp, M);
    SEQM_CHECK_LAUNCH("hcore_diag_kernel");
  }
  prof_mark("end", st);
  return SEQM_OK;
}

int seqm_w_expand(const seqm_topology_t* topo, seqm_w_t w, double* w_dense, void* stream) {
  SEQM_REQUIRE(topo && w_dense, "seqm_w_expand: null pointer");
  if (topo->npairs == 0) return SEQM_OK;
  w_expand_kernel<<<cdiv(topo->npairs * 100, 256), 256, 0, (cudaStream_t)stream>>>(*topo, w, w_dense);
  SEQM_CHECK_LAUNCH("w_expand_kernel");
  return SEQM_OK;
}

int seqm_pair_grad(const seqm_topology_t* topo, const void* atompar, const double* xij, const double* rij,
                   int ngauss, const double* P, const double* D, const double* gE, const double* gNuc,
                   double* grad_xij, double* grad_rij, void* stream) {
  SEQM_REQUIRE(topo && atompar && xij && rij && P && grad_xij && grad_rij, "seqm_pair_grad: null pointer");
  const seqm_topology_t& t = *topo;
  cudaStream_t st = (cudaStream_t)stream;
  const AtomPar* ap = (const AtomPar*)atompar;
  prof_mark("pair_grad", st);
  const long long mXX = cdiv(t.ncls[2], PGX_NT), mXH = cdiv(t.ncls[1], PGX_NT), mHH = cdiv(t.ncls[0], PGX_NT);
  const int nclasses = (mXX > 0) + (mXH > 0) + (mHH > 0);
  if (g_pair_grad_reverse && nclasses > 1 && mXX + mXH + mHH <= PAIR_MERGE_WAVES * 148 * 8) {      // 8 one-warp CTAs per SM
    pair_grad_all_kernel<<<(unsigned)(mXX + mXH + mHH), PGX_NT, 0, st>>>(t, ap, xij, rij, ngauss, P, D, gE, gNuc, grad_xij,
                                                                       grad_rij, (int)mXX, (int)mXH);
    SEQM_CHECK_LAUNCH("pair_grad_all_kernel");
    prof_mark("end", st);
    return SEQM_OK;
  }
  if (t.ncls[2] > 0 && g_pair_grad_reverse) {
    pair_grad_xx_kernel<<<cdiv(t.ncls[2], PGX_NT), PGX_NT, 0, st>>>(t, ap, xij, rij, ngauss, P, D, gE, gNuc, grad_xij, grad_rij);
    SEQM_CHECK_LAUNCH("pair_grad_xx_kernel");
  } else if (t.ncls[2] > 0) {
    pair_grad_kernel<SEQM_CLS_XX><<<cdiv(t.ncls[2], 128), 128, 0, st>>>(t, ap, xij, rij, ngauss, P, D, gE, gNuc, grad_xij, grad_rij);
    SEQM_CHECK_LAUNCH("pair_grad_kernel<XX>");
  }
  if (t.ncls[1] > 0) {
    pair_grad_kernel<SEQM_CLS_XH><<<cdiv(t.ncls[1], 128), 128, 0, st>>>(t, ap, xij, rij, ngauss, P, D, gE, gNuc, grad_xij, grad_rij);
    SEQM_CHECK_LAUNCH("pair_grad_kernel<XH>");
  }
  if (t.ncls[0] > 0) {
    pair_grad_kernel<SEQM_CLS_HH><<<cdiv(t.ncls[0], 128), 128, 0, st>>>(t, ap, xij, rij, ngauss, P, D, gE, gNuc, grad_xij, grad_rij);
    SEQM_CHECK_LAUNCH("pair_grad_kernel<HH>");
  }
  prof_mark("end", st);
  return SEQM_OK;
}

void seqm_pair_grad_reverse(int on) { g_pair_grad_reverse = on; }

int seqm_pair_geometry(const seqm_topology_t* topo, const double* R, double lcf, double* xij, double* rij,
                       void* stream) {
  SEQM_REQUIRE(topo && R && xij && rij, "seqm_pair_geometry: null pointer");
  if (topo->npairs == 0) return SEQM_OK;
  prof_mark("pair_geometry", (cudaStream_t)stream);
  pair_geometry_kernel<<<cdiv(topo->npairs, 256), 256, 0, (cudaStream_t)stream>>>(*topo, R, lcf, xij, rij);
  SEQM_CHECK_LAUNCH("pair_geometry_kernel");
  prof_mark("end", (cudaStream_t)stream);
  return SEQM_OK;
}

int seqm_pair_geometry_grad(const seqm_topology_t* topo, const double* xij, const double* rij, const double* gx,
                            const double* gr, double lcf, double* gR, void* stream) {
  SEQM_REQUIRE(topo && xij && rij && gx && gr && gR, "seqm_pair_geometry_grad: null pointer");
  if (topo->natoms == 0) return SEQM_OK;
  prof_mark("pair_geometry_grad", (cudaStream_t)stream);
  pair_geometry_grad_kernel<<<cdiv(topo->natoms * 32, 256), 256, 0, (cudaStream_t)stream>>>(*topo, xij, rij, gx, gr,
                                                                                          lcf, gR);
  SEQM_CHECK_LAUNCH("pair_geometry_grad_kernel");
  prof_mark("end", (cudaStream_t)stream);
  return SEQM_OK;
}

}  // extern "C"
