// device code of scf_small_kernels.cu (one warp per molecule)
#pragma once
#include "eigh_warp.cuh"

namespace seqm {

#define KM(k) ((k) == 0 ? 0 : ((k) <= 2 ? 1 : ((k) <= 5 ? 2 : 3)))
#define KN(k) ((k) - KM(k) * (KM(k) + 1) / 2)
#define WT(k) ((KM(k) == KN(k)) ? 1.0 : 2.0)

constexpr int SNFOCK = 5;
enum SmallMix { SMIX_DIRECT = 0, SMIX_CONST = 1, SMIX_ADAPTIVE = 2, SMIX_TAKE = 3 };

struct SmallScfArgs {
  seqm_topology_t t;
  seqm_w_t w;
  const double *M, *gss, *gpp, *gsp, *gp2, *hsp;
  const int32_t *nheavy, *nhydro, *nocc;
  const int32_t* mol_atom0;      // [B+1] first real atom of every molecule
  double *P, *F, *E;             // P in/out (padded), F out (padded, zero filled by the caller), E out
  uint8_t* nc;
  int32_t* niter;
  double *pairvec, *Pold, *Hpk, *FOCK, *FPPF, *EMAT, *Vstore;   // Hpk: packed core Hamiltonian scratch [B][nn]
  unsigned long long* stat;      // [0] molecules left unconverged, [1] largest iteration count
  size_t nn;                     // doubles per molecule in Pold / one ring slot
  int converger, max_iter, check_degeneracy;
  double alpha, eps, tol;
};

// exchange block and Coulomb vectors of one pair (fock.py:76-139); P and F packed in shared memory
// (p = pair index of class slot `slot`, si / sj = the slots of its atoms inside the molecule: three levels of dependent
// index loads that the caller issues for all class passes at once)
template <int CLS, int LD>
__device__ __forceinline__ void small_fock_pair(const SmallScfArgs& A, int mol, int slot, int p, int si, int sj, int nh,
                                                int nh4, const double* __restrict__ Ps, double* __restrict__ Fs) {
  const seqm_topology_t& t = A.t;
  const int S = (int)t.molsize;
  const int oi = si < nh ? 4 * si : nh4 + si - nh, oj = sj < nh ? 4 * sj : nh4 + sj - nh;
  const double* Mb = A.M + (((size_t)mol * S + si) * S + sj) * 16;
  double* pv = A.pairvec + (size_t)p * 20;
  constexpr int NA = (CLS == SEQM_CLS_HH) ? 1 : 10;
  constexpr int NB = (CLS == SEQM_CLS_XX) ? 10 : 1;
  constexpr int OA = (CLS == SEQM_CLS_HH) ? 1 : 4;
  constexpr int OB = (CLS == SEQM_CLS_XX) ? 4 : 1;
  double pA[NA], pB[NB];
  {
    const double* a = Ps + oi * LD + oi;
    const double* b = Ps + oj * LD + oj;
#pragma unroll
    for (int k = 0; k < NA; ++k) pA[k] = WT(k) * a[KN(k) * LD + KM(k)];
#pragma unroll
    for (int l = 0; l < NB; ++l) pB[l] = WT(l) * b[KN(l) * LD + KM(l)];
  }
  double Pab[OA][OB];
#pragma unroll
  for (int mu = 0; mu < OA; ++mu)
#pragma unroll
    for (int nu = 0; nu < OB; ++nu) Pab[mu][nu] = Ps[(oi + mu) * LD + oj + nu];
  double toA[NA], toB[NB], Fab[OA][OB];
#pragma unroll
  for (int k = 0; k < NA; ++k) toA[k] = 0.0;
#pragma unroll
  for (int l = 0; l < NB; ++l) toB[l] = 0.0;
#pragma unroll
  for (int mu = 0; mu < OA; ++mu)
#pragma unroll
    for (int nu = 0; nu < OB; ++nu) Fab[mu][nu] = 0.0;
  const double* wc = CLS == SEQM_CLS_HH ? A.w.wHH : (CLS == SEQM_CLS_XH ? A.w.wXH : A.w.wXX);
  const double* wp = (CLS == SEQM_CLS_HH) ? wc + slot : wc + ((size_t)(slot >> 5) * (NA * NB)) * 32 + (slot & 31);
#pragma unroll
  for (int k = 0; k < NA; ++k) {
#pragma unroll
    for (int l = 0; l < NB; ++l) {
      const double wv = (CLS == SEQM_CLS_HH) ? wp[0] : wp[(size_t)(k * NB + l) * 32];
      toA[k] += wv * pB[l];
      toB[l] += pA[k] * wv;
      const int m = KM(k), n = KN(k), a = KM(l), b = KN(l);
      Fab[m][a] += Pab[n][b] * wv;
      if (a != b) Fab[m][b] += Pab[n][a] * wv;
      if (m != n) {
        Fab[n][a] += Pab[m][b] * wv;
        if (a != b) Fab[n][b] += Pab[m][a] * wv;
      }
    }
  }
#pragma unroll
  for (int k = 0; k < NA; ++k) pv[k] = toA[k];
#pragma unroll
  for (int l = 0; l < NB; ++l) pv[10 + l] = toB[l];
#pragma unroll
  for (int mu = 0; mu < OA; ++mu)
#pragma unroll
    for (int nu = 0; nu < OB; ++nu) {
      const double f = Mb[mu * 4 + nu] - 0.5 * Fab[mu][nu];
      Fs[(oi + mu) * LD + oj + nu] = f;
      Fs[(oj + nu) * LD + oi + mu] = f;
    }
}

}  // namespace seqm
