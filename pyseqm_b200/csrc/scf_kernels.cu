// K3c: the SCF driver.  scf_forward0 (constant mixing), scf_forward1 (adaptive mixing, MOPAC cnvg) and
// scf_forward2 (2 direct + 1 adaptive + Pulay/DIIS with a ring of 5 Fock matrices) of the reference, with
// every per-molecule update masked by the device-resident `notconverged` flags exactly as the reference
// masks with notconverged[...] — converged molecules are frozen, CTAs assigned to them exit at once.
// The host loop only reads back one integer (how many molecules are still unconverged) per iteration,
// the analogue of the reference's notconverged.any().
// reference: seqm/seqm_functions/scf_loop.py:32-510 (iteration schedule and EMAT/EVEC ring arithmetic,
// including the stale-lower-triangle quirk after the ring wraps, scf_loop.py:460-490)
#include <cstring>
#include <vector>

#include "abi_common.cuh"
#include "dgemm.cuh"

namespace seqm {

constexpr int NFOCK = 5;
int g_scf_fused = 1;      // 0: always the host-driven loop (test hook)

enum MixMode { MIX_DIRECT = 0, MIX_CONST = 1, MIX_ADAPTIVE = 2, MIX_TAKE = 3 };

// DIRECT: Pold = P, P = Pnew.  CONST: P = alpha P + (1-alpha) Pnew.  ADAPTIVE: fac from the padded diagonals,
// Pold = P, P = (1+fac) Pnew - fac P.  TAKE: P = Pnew.  All per-molecule kernels below are cut into chunks of
// RED_CHUNK matrix elements, one CTA per (molecule, chunk): blockIdx.x = molecule * chunks + chunk.
__global__ void __launch_bounds__(256)
mix_fac_kernel(int n4, const uint8_t* __restrict__ active, const double* __restrict__ Pnew,
               const double* __restrict__ P, const double* __restrict__ Pold, double* __restrict__ fac_out,
               const int32_t* __restrict__ nheavy, const int32_t* __restrict__ nhydro) {
  const int mol = blockIdx.x;
  if (active && !active[mol]) return;
  const size_t base = (size_t)mol * n4 * n4;
  const int nh4 = 4 * nheavy[mol], m = nh4 + nhydro[mol];      // the diagonal elements that belong to an orbital (the others are zero)
  __shared__ double red[2][8];
  double num = 0.0, den = 0.0;
  for (int ip = threadIdx.x; ip < m; ip += blockDim.x) {
    const int i = ip < nh4 ? ip : nh4 + 4 * (ip - nh4);
    const size_t d = base + (size_t)i * n4 + i;
    const double dn = Pnew[d], dc = P[d], dold = Pold[d];
    num += (dn - dc) * (dn - dc);
    const double u = dn - dc * 2.0 + dold;
    den += u * u;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    num += __shfl_xor_sync(0xffffffffu, num, o);
    den += __shfl_xor_sync(0xffffffffu, den, o);
  }
  if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = num; red[1][threadIdx.x >> 5] = den; }
  __syncthreads();
  if (threadIdx.x == 0) {
    num = 0.0; den = 0.0;
    for (int q = 0; q < (int)(blockDim.x >> 5); ++q) { num += red[0][q]; den += red[1][q]; }
    fac_out[mol] = sqrt(num / den);
  }
}

__global__ void __launch_bounds__(256)
mix_kernel(int n4, int chunks, int mode, double alpha, const uint8_t* __restrict__ active,
           const double* __restrict__ Pnew, double* __restrict__ P, double* __restrict__ Pold,
           const double* __restrict__ facs, const int32_t* __restrict__ nheavy, const int32_t* __restrict__ nhydro) {
  const int mol = blockIdx.x / chunks, chunk = blockIdx.x - mol * chunks;
  if (active && !active[mol]) return;
  // only the rows that hold an orbital: the density step writes nothing else, and nothing reads the other rows of Pold
  const int nh4 = 4 * nheavy[mol], m = nh4 + nhydro[mol];
  const size_t nn = (size_t)m * n4, base = (size_t)mol * n4 * n4;
  const size_t lo = (size_t)chunk * RED_CHUNK, hi = lo + RED_CHUNK < nn ? lo + RED_CHUNK : nn;
  const double fac = mode == MIX_ADAPTIVE ? facs[mol] : 0.0;
  for (size_t idx0 = lo + threadIdx.x; idx0 < hi; idx0 += blockDim.x) {
    int ip, jp;
    row_col(idx0, n4, ip, jp);
    const size_t idx = (size_t)(ip < nh4 ? ip : nh4 + 4 * (ip - nh4)) * n4 + jp;
    const double pn = Pnew[base + idx];
    const double pc = P[base + idx];
    if (mode == MIX_DIRECT) { Pold[base + idx] = pc; P[base + idx] = pn; }
    else if (mode == MIX_CONST) { P[base + idx] = alpha * pc + (1.0 - alpha) * pn; }
    else if (mode == MIX_ADAPTIVE) { Pold[base + idx] = pc; P[base + idx] = (1.0 + fac) * pn - fac * pc; }
    else { P[base + idx] = pn; }
  }
}

// err = |Enew - E|, E = Enew, niter++, notconverged = err > eps (for the still active molecules);
// counts the molecules left and tracks max err (as an ordered integer image of the non-negative double)
__global__ void finish_step_kernel(int nmol, double eps, const double* __restrict__ Enew, double* __restrict__ E,
                                   double* __restrict__ err, uint8_t* __restrict__ nc, int32_t* __restrict__ niter,
                                   unsigned long long* __restrict__ stat) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nmol) return;
  if (nc[b]) {
    const double en = Enew[b];
    err[b] = fabs(en - E[b]);
    E[b] = en;
    niter[b] += 1;
  }
  const double e = err[b];
  const bool still = e > eps;
  nc[b] = still ? 1 : 0;
  if (still) atomicAdd(&stat[0], 1ULL);
  atomicMax(&stat[1], (unsigned long long)__double_as_longlong(e));
}

// packed copies of F and P for the Pulay commutator of hydrogen-rich systems: Fp[b][i][j] = F[b][pad(i)][pad(j)]
// (heavy atoms first, 4 orbitals each, then one orbital per hydrogen; pack.py:7-13), zero beyond the packed size
__global__ void __launch_bounds__(256)
pack_pair_kernel(int n4, int mld, int chunks, const uint8_t* __restrict__ active, const int32_t* __restrict__ nheavy,
                 const int32_t* __restrict__ nhydro, const double* __restrict__ F, const double* __restrict__ P,
                 double* __restrict__ Fp, double* __restrict__ Pp) {
  const int mol = blockIdx.x / chunks, chunk = blockIdx.x - mol * chunks;
  if (active && !active[mol]) return;
  const int nh4 = 4 * nheavy[mol], m = nh4 + nhydro[mol];
  const size_t tot = (size_t)mld * mld, src = (size_t)mol * n4 * n4, dst = (size_t)mol * tot;
  const size_t lo = (size_t)chunk * RED_CHUNK, hi = lo + RED_CHUNK < tot ? lo + RED_CHUNK : tot;
  for (size_t idx = lo + threadIdx.x; idx < hi; idx += blockDim.x) {
    int i, j;
    row_col(idx, mld, i, j);
    double f = 0.0, q = 0.0;
    if (i < m && j < m) {
      const int pi = i < nh4 ? i : nh4 + 4 * (i - nh4), pj = j < nh4 ? j : nh4 + 4 * (j - nh4);
      f = F[src + (size_t)pi * n4 + pj];
      q = P[src + (size_t)pi * n4 + pj];
    }
    Fp[dst + idx] = f;
    Pp[dst + idx] = q;
  }
}

// Pulay store (padded ring): FOCK[c] = F, FPPF[c] = triu(G - G^T) with G = F P
__global__ void __launch_bounds__(256)
pulay_store_kernel(int n4, int chunks, int slot, const uint8_t* __restrict__ active, const double* __restrict__ F,
                   const double* __restrict__ G, double* __restrict__ FOCK, double* __restrict__ FPPF) {
  const int mol = blockIdx.x / chunks, chunk = blockIdx.x - mol * chunks;
  if (active && !active[mol]) return;
  const size_t nn = (size_t)n4 * n4;
  const size_t src = (size_t)mol * nn;
  const size_t dst = ((size_t)mol * NFOCK + slot) * nn;
  const size_t lo = (size_t)chunk * RED_CHUNK, hi = lo + RED_CHUNK < nn ? lo + RED_CHUNK : nn;
  for (size_t idx = lo + threadIdx.x; idx < hi; idx += blockDim.x) {
    int i, j;
    row_col(idx, n4, i, j);
    FOCK[dst + idx] = F[src + idx];
    FPPF[dst + idx] = (j >= i) ? G[src + idx] - G[src + (size_t)j * n4 + i] : 0.0;
  }
}

// partial sums of EMAT[c][s] = <FPPF[c], FPPF[s]> for s < cF: blockIdx.x = (molecule * NFOCK + s) * chunks + chunk
// (items with s >= cF are not launched: grid covers s < cF through the index arithmetic below)
__global__ void __launch_bounds__(256)
pulay_emat_part_kernel(size_t nn, int chunks, int cF, int slot, const uint8_t* __restrict__ active,
                       const double* __restrict__ FPPF, double* __restrict__ part) {
  // nn = doubles per ring slot (n4^2, or mld^2 for the packed ring)
  const int item = blockIdx.x / chunks, chunk = blockIdx.x - item * chunks;
  const int mol = item / cF, s = item - mol * cF;
  if (active && !active[mol]) return;
  const double* a = FPPF + ((size_t)mol * NFOCK + slot) * nn;
  const double* b = FPPF + ((size_t)mol * NFOCK + s) * nn;
  const size_t lo = (size_t)chunk * RED_CHUNK, hi = lo + RED_CHUNK < nn ? lo + RED_CHUNK : nn;
  double acc = 0.0;
  for (size_t idx = lo + threadIdx.x; idx < hi; idx += blockDim.x) acc += a[idx] * b[idx];
  __shared__ double red[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int q = 0; q < (int)(blockDim.x >> 5); ++q) t += red[q];
    part[(size_t)item * chunks + chunk] = t;
  }
}

// where item (molecule, s) of the partial sums lands in EMAT: EMAT[mol][slot][s]
__global__ void pulay_emat_index_kernel(int nitems, int cF, int slot, int32_t* __restrict__ index) {
  const int item = blockIdx.x * blockDim.x + threadIdx.x;
  if (item >= nitems) return;
  const int mol = item / cF, s = item - mol * cF;
  index[item] = mol * 36 + slot * 6 + s;
}

// EVEC = EMAT + tril(EMAT,-1)^T; EVEC[:cF,:cF] /= EVEC[c][c]; coeff = -inverse(EVEC[:cF+1,:cF+1])[:-1,-1]
__global__ void pulay_coeff_kernel(int nmol, int slot, int cF, const uint8_t* __restrict__ active,
                                   const double* __restrict__ EMAT, double* __restrict__ coeff) {
  const int mol = blockIdx.x * blockDim.x + threadIdx.x;
  if (mol >= nmol) return;
  if (active && !active[mol]) return;
  {
    const double* em = EMAT + (size_t)mol * 36;
    const int d = cF + 1;
    double a[6][7];
    for (int i = 0; i < 6; ++i)
      for (int j = 0; j < 6; ++j) a[i][j] = em[i * 6 + j] + (j > i ? em[j * 6 + i] : 0.0);
    const double sc = a[slot][slot];
    for (int i = 0; i < cF; ++i)
      for (int j = 0; j < cF; ++j) a[i][j] = a[i][j] / sc;
    // solve a[:d,:d] x = e_{d-1} by Gaussian elimination with partial pivoting: x = last column of the inverse
    for (int i = 0; i < d; ++i) a[i][6] = (i == d - 1) ? 1.0 : 0.0;
    for (int c = 0; c < d; ++c) {
      int piv = c;
      double best = fabs(a[c][c]);
      for (int r = c + 1; r < d; ++r)
        if (fabs(a[r][c]) > best) { best = fabs(a[r][c]); piv = r; }
      if (piv != c) {
        for (int j = 0; j < 7; ++j) { const double tmp = a[c][j]; a[c][j] = a[piv][j]; a[piv][j] = tmp; }
      }
      const double inv = 1.0 / a[c][c];
      for (int r = c + 1; r < d; ++r) {
        const double f = a[r][c] * inv;
        for (int j = c; j < d; ++j) a[r][j] -= f * a[c][j];
        a[r][6] -= f * a[c][6];
      }
    }
    double x[6];
    for (int r = d - 1; r >= 0; --r) {
      double sum = a[r][6];
      for (int j = r + 1; j < d; ++j) sum -= a[r][j] * x[j];
      x[r] = sum / a[r][r];
    }
    for (int s = 0; s < cF; ++s) coeff[(size_t)mol * NFOCK + s] = -x[s];
  }
}

// F = sum_s coeff_s FOCK[s]
__global__ void __launch_bounds__(256)
pulay_extrapolate_kernel(int n4, int chunks, int cF, const uint8_t* __restrict__ active,
                         const double* __restrict__ coeffs, const double* __restrict__ FOCK, double* __restrict__ F) {
  const int mol = blockIdx.x / chunks, chunk = blockIdx.x - mol * chunks;
  if (active && !active[mol]) return;
  double coeff[NFOCK];
#pragma unroll
  for (int s = 0; s < NFOCK; ++s) coeff[s] = s < cF ? coeffs[(size_t)mol * NFOCK + s] : 0.0;
  const size_t nn = (size_t)n4 * n4;
  const double* fo = FOCK + (size_t)mol * NFOCK * nn;
  double* out = F + (size_t)mol * nn;
  const size_t lo = (size_t)chunk * RED_CHUNK, hi = lo + RED_CHUNK < nn ? lo + RED_CHUNK : nn;
  for (size_t idx = lo + threadIdx.x; idx < hi; idx += blockDim.x) {
    double acc = 0.0;
#pragma unroll
    for (int s = 0; s < NFOCK; ++s)
      if (s < cF) acc += fo[(size_t)s * nn + idx] * coeff[s];
    out[idx] = acc;
  }
}

// Packed Pulay ring (hydrogen-rich systems, see packed_comm in seqm_scf): a ring slot holds the PACKED matrices
// (mld x mld instead of n4 x n4, 4-5 x fewer bytes for organic molecules): FOCK[c] = Fp, FPPF[c] = triu(Gp - Gp^T).
// The packing keeps the order of the orbitals, so these are the non-zero elements of the padded matrices in the same
// row-major order; only the grouping of the EMAT partial sums (chunks of the packed instead of the padded matrix) differs.
__global__ void __launch_bounds__(256)
pulay_store_packed_kernel(int mld, int chunks, int slot, const uint8_t* __restrict__ active, const double* __restrict__ Fp,
                          const double* __restrict__ Gp, double* __restrict__ FOCK, double* __restrict__ FPPF) {
  const int mol = blockIdx.x / chunks, chunk = blockIdx.x - mol * chunks;
  if (active && !active[mol]) return;
  const size_t pk = (size_t)mld * mld;
  const size_t src = (size_t)mol * pk, dst = ((size_t)mol * NFOCK + slot) * pk;
  const size_t lo = (size_t)chunk * RED_CHUNK, hi = lo + RED_CHUNK < pk ? lo + RED_CHUNK : pk;
  for (size_t idx = lo + threadIdx.x; idx < hi; idx += blockDim.x) {
    int i, j;
    row_col(idx, mld, i, j);
    FOCK[dst + idx] = Fp[src + idx];
    FPPF[dst + idx] = (j >= i) ? Gp[src + idx] - Gp[src + (size_t)j * mld + i] : 0.0;
  }
}

// F = sum_s coeff_s FOCK[s] from the packed ring, written to the orbital positions of the padded F (the rest of F is
// zero and stays zero: the Fock build writes orbital positions only)
__global__ void __launch_bounds__(256)
pulay_extrapolate_packed_kernel(int n4, int mld, int chunks, int cF, const uint8_t* __restrict__ active,
                                const int32_t* __restrict__ nheavy, const int32_t* __restrict__ nhydro,
                                const double* __restrict__ coeffs, const double* __restrict__ FOCK, double* __restrict__ F) {
  const int mol = blockIdx.x / chunks, chunk = blockIdx.x - mol * chunks;
  if (active && !active[mol]) return;
  double coeff[NFOCK];
#pragma unroll
  for (int s = 0; s < NFOCK; ++s) coeff[s] = s < cF ? coeffs[(size_t)mol * NFOCK + s] : 0.0;
  const int nh4 = 4 * nheavy[mol], m = nh4 + nhydro[mol];
  const size_t pk = (size_t)mld * mld;
  const double* fo = FOCK + (size_t)mol * NFOCK * pk;
  double* out = F + (size_t)mol * n4 * n4;
  const size_t lo = (size_t)chunk * RED_CHUNK, hi = lo + RED_CHUNK < pk ? lo + RED_CHUNK : pk;
  for (size_t idx = lo + threadIdx.x; idx < hi; idx += blockDim.x) {
    int i, j;
    row_col(idx, mld, i, j);
    if (i >= m || j >= m) continue;
    double acc = 0.0;
#pragma unroll
    for (int s = 0; s < NFOCK; ++s)
      if (s < cF) acc += fo[(size_t)s * pk + idx] * coeff[s];
    const int pi = i < nh4 ? i : nh4 + 4 * (i - nh4), pj = j < nh4 ? j : nh4 + 4 * (j - nh4);
    out[(size_t)pi * n4 + pj] = acc;
  }
}

__global__ void init_flags_kernel(int nmol, uint8_t* nc, double* err, int32_t* niter, double* EMAT) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nmol) return;
  nc[b] = 1;
  err[b] = 1.0;
  niter[b] = 0;
  if (EMAT) {
    for (int i = 0; i < 6; ++i)
      for (int j = 0; j < 6; ++j) EMAT[(size_t)b * 36 + i * 6 + j] = (i > j) ? -1.0 : 0.0;
  }
}

struct ScfWork {
  double *Pnew, *Pold, *Enew, *err, *G, *FOCK, *FPPF, *EMAT, *Vstore;
  double *red, *fac, *coeff;      // chunked-reduction partial sums [B][NFOCK][chunks], mixing factors [B], DIIS coefficients [B][NFOCK]
  int32_t* emat_index;
  unsigned long long* stat;
  int32_t* mol_atom0;
  char* sp2;
  size_t nn, total;      // nn = doubles per molecule in one matrix slab
};

static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

static ScfWork carve(char* base, int64_t nmol, int64_t molsize, const seqm_scf_options_t* opt) {
  ScfWork w{};
  const size_t nn = (size_t)16 * molsize * molsize > 64 ? (size_t)16 * molsize * molsize : 64;   // >= 8 x 8 (K3s)
  w.nn = nn;
  size_t off = 0;
  auto take = [&](size_t bytes) { char* p = base ? base + off : nullptr; off += align256(bytes); return p; };
  w.Pnew = (double*)take(nmol * nn * 8);
  w.Pold = (double*)take(nmol * nn * 8);
  w.Enew = (double*)take(nmol * 8);
  w.err = (double*)take(nmol * 8);
  w.stat = (unsigned long long*)take(64);
  w.mol_atom0 = (int32_t*)take(scf_small_scratch_bytes(nmol));
  w.red = (double*)take((size_t)nmol * NFOCK * red_chunks(nn) * 8);
  w.fac = (double*)take(nmol * 8);
  w.coeff = (double*)take(nmol * NFOCK * 8);
  w.emat_index = (int32_t*)take(nmol * NFOCK * 4);
  if (opt->converger == 2) {
    w.G = (double*)take(nmol * nn * 8);
    w.FOCK = (double*)take(nmol * NFOCK * nn * 8);
    w.FPPF = (double*)take(nmol * NFOCK * nn * 8);
    w.EMAT = (double*)take(nmol * 36 * 8);
  }
  if (opt->use_sp2)
    w.sp2 = take(g_scf_sp2_nranks >= 1 ? seqm_sp2_nccl_workspace(nmol, molsize, g_scf_sp2_nranks)
                                       : seqm_sp2_workspace(nmol, molsize));
  else w.Vstore = (double*)take((size_t)nmol * (16 * molsize * molsize > 1024 ? 16 * molsize * molsize : 1024) * 8);
  w.total = off;
  return w;
}

}  // namespace seqm

using namespace seqm;

extern "C" {

void seqm_scf_fused(int on) { g_scf_fused = on; }

size_t seqm_scf_workspace(int64_t nmol, int64_t molsize, const seqm_scf_options_t* opt) {
  if (!opt) return 0;
  return carve(nullptr, nmol, molsize, opt).total;
}

int seqm_scf(const seqm_topology_t* topo, seqm_w_t w, const double* M, const double* g_ss, const double* g_pp,
             const double* g_sp, const double* g_p2, const double* h_sp, const int32_t* nheavy,
             const int32_t* nhydro, const int32_t* nocc, int mmax, const seqm_scf_options_t* opt, double* P,
             double* F, double* Eelec, uint8_t* notconverged, int32_t* niter, double* pairvec, void* work,
             int32_t* iterations_out, void* stream) {
  SEQM_REQUIRE(topo && M && g_ss && g_pp && g_sp && g_p2 && h_sp && nheavy && nhydro && nocc && opt && P && F &&
                   Eelec && notconverged && niter && pairvec && work,
               "seqm_scf: null pointer");
  SEQM_REQUIRE(opt->converger >= 0 && opt->converger <= 2, "seqm_scf: scf_converger must be [0, alpha], [1] or [2]");
  const seqm_topology_t& t = *topo;
  cudaStream_t st = (cudaStream_t)stream;
  SEQM_REQUIRE(t.molsize > 0 && t.molsize < 16384, "seqm_scf: molsize out of range (a padded matrix must have fewer than 2^32 elements)");
  const int B = (int)t.nmol, S = (int)t.molsize, n4 = 4 * S;
  const size_t nn = (size_t)n4 * n4;
  if (B == 0) { if (iterations_out) *iterations_out = 0; return SEQM_OK; }
  ScfWork wk = carve((char*)work, B, S, opt);
  uint8_t* nc = notconverged;
  int rc;
  const int mp_small = ((mmax < 1 ? 1 : mmax) + 1) & ~1;
  if (!opt->use_sp2 && !opt->debug && mp_small <= 32 && g_scf_fused) {
    // K3s: the whole loop in one launch, one warp per molecule
    const int max_iter_s = opt->max_iter > 0 ? opt->max_iter : 1000;
    SEQM_CUDA(cudaMemsetAsync(F, 0, (size_t)B * nn * 8, st));
    SEQM_CUDA(cudaMemsetAsync(wk.stat, 0, 16, st));
    rc = launch_scf_small(t, w, M, g_ss, g_pp, g_sp, g_p2, h_sp, nheavy, nhydro, nocc, mp_small, opt->converger,
                          opt->alpha, opt->eps, max_iter_s, eigh_tolerance(), P, F, Eelec, nc, niter, pairvec, wk.Pold,
                          wk.Pnew, wk.FOCK, wk.FPPF, wk.EMAT, wk.Vstore, wk.stat, wk.mol_atom0, wk.nn, st);
    if (rc) return rc;
    if (opt->async_status) {
      SEQM_REQUIRE(iterations_out, "seqm_scf: async_status needs the page-locked status buffer");
      SEQM_CUDA(cudaMemcpyAsync(iterations_out, wk.stat, 16, cudaMemcpyDeviceToHost, st));
      return SEQM_OK;
    }
    unsigned long long hs[2] = {0, 0};
    SEQM_CUDA(cudaMemcpyAsync(hs, wk.stat, 16, cudaMemcpyDeviceToHost, st));
    SEQM_CUDA(cudaStreamSynchronize(st));
    if (iterations_out) { iterations_out[0] = (int32_t)hs[1]; iterations_out[1] = (int32_t)hs[0]; }
    return SEQM_OK;
  }

  prof_mark("scf_setup", st);
  SEQM_CUDA(cudaMemsetAsync(wk.Pnew, 0, (size_t)B * nn * 8, st));
  SEQM_CUDA(cudaMemsetAsync(wk.Pold, 0, (size_t)B * nn * 8, st));
  SEQM_CUDA(cudaMemsetAsync(F, 0, (size_t)B * nn * 8, st));
  init_flags_kernel<<<cdiv(B, 256), 256, 0, st>>>(B, nc, wk.err, niter, wk.EMAT);
  SEQM_CHECK_LAUNCH("init_flags_kernel");
  // F = fock(P), Eelec
  rc = launch_fock(t, w, P, M, g_ss, g_pp, g_sp, g_p2, h_sp, nullptr, F, Eelec, pairvec, wk.red, st, nheavy, nhydro);
  if (rc) return rc;

  unsigned long long* hstat = host_status_slot();
  SEQM_REQUIRE(hstat, "seqm_scf: no page-locked status slot");
  int k = 0;
  long long left = B;
  const int max_iter = opt->max_iter > 0 ? opt->max_iter : 1000;

  int ncalls_density = 0;
  auto density = [&](double* out) -> int {
    if (opt->use_sp2 && g_scf_sp2_nranks >= 1)      // tile-sharded over the ranks of seqm_scf_sp2_comm()
      return seqm_sp2_density_nccl(g_scf_sp2_comm, g_scf_sp2_nranks, B, S, mmax, F, nheavy, nhydro, nocc, opt->sp2_eps,
                                   nc, out, nullptr, wk.sp2, st);
    if (opt->use_sp2)
      return seqm_sp2_density(B, S, mmax, F, nheavy, nhydro, nocc, opt->sp2_eps, nc, out, nullptr, wk.sp2, st);
    const int warm = ncalls_density > 0 ? 1 : 0;
    ++ncalls_density;
    return launch_eigh_density(B, S, mmax, F, nheavy, nhydro, nocc, nc, out, nullptr, nullptr, nullptr, wk.Vstore, warm, st);
  };
  const int chunks = red_chunks(nn);
  SEQM_REQUIRE((size_t)B * NFOCK * chunks < ((size_t)1 << 31), "seqm_scf: batch too large for the chunked kernels");
  // FOCK[slot] = F, FPPF[slot] = triu(FP - PF) (G = F P given), EMAT[slot][s] = <FPPF[slot], FPPF[s]>, s < cF
  // hydrogen-rich systems (3 packed matrices fit into the one padded scratch matrix): the product F P of the Pulay
  // commutator runs on the packed matrices - (n4 / m)^3 fewer flops, 8 x for water
  const int mld = (mmax + 3) & ~3;
  const bool packed_comm = (size_t)3 * mld * mld <= nn;
  auto pulay_store = [&](int slot, int cF_) -> int {
    if (packed_comm) {
      const size_t pk = (size_t)mld * mld;
      double *Fp = wk.G, *Pp = wk.G + (size_t)B * pk, *Gp = wk.G + (size_t)2 * B * pk;
      const int pchunks = red_chunks(pk);
      prof_mark("pulay_store", st);
      pack_pair_kernel<<<(unsigned)((size_t)B * pchunks), 256, 0, st>>>(n4, mld, pchunks, nc, nheavy, nhydro, F, P, Fp, Pp);
      SEQM_CHECK_LAUNCH("pack_pair_kernel");
      int r_ = launch_dgemm_batched(mld, mld, Fp, Pp, Gp, (long long)pk, B, nc, nullptr, st);
      if (r_) return r_;
      prof_mark("pulay_store", st);
      pulay_store_packed_kernel<<<(unsigned)((size_t)B * pchunks), 256, 0, st>>>(mld, pchunks, slot, nc, Fp, Gp, wk.FOCK, wk.FPPF);
      SEQM_CHECK_LAUNCH("pulay_store_packed_kernel");
    } else {
      int r_ = launch_dgemm_batched(n4, n4, F, P, wk.G, (long long)nn, B, nc, nullptr, st);
      if (r_) return r_;
      prof_mark("pulay_store", st);
      pulay_store_kernel<<<(unsigned)((size_t)B * chunks), 256, 0, st>>>(n4, chunks, slot, nc, F, wk.G, wk.FOCK, wk.FPPF);
      SEQM_CHECK_LAUNCH("pulay_store_kernel");
    }
    const size_t rlen = packed_comm ? (size_t)mld * mld : nn;      // doubles per ring slot
    const int rchunks = packed_comm ? red_chunks(rlen) : chunks;
    pulay_emat_part_kernel<<<(unsigned)((size_t)B * cF_ * rchunks), 256, 0, st>>>(rlen, rchunks, cF_, slot, nc, wk.FPPF, wk.red);
    SEQM_CHECK_LAUNCH("pulay_emat_part_kernel");
    pulay_emat_index_kernel<<<cdiv((long long)B * cF_, 256), 256, 0, st>>>(B * cF_, cF_, slot, wk.emat_index);
    SEQM_CHECK_LAUNCH("pulay_emat_index_kernel");
    return launch_chunk_sum(B * cF_, rchunks, nc, cF_, wk.red, wk.EMAT, wk.emat_index, st);
  };
  auto refock_and_finish = [&](bool store, int& counter, int& cF) -> int {
    int r = launch_fock(t, w, P, M, g_ss, g_pp, g_sp, g_p2, h_sp, nc, F, wk.Enew, pairvec, wk.red, st, nheavy, nhydro);
    if (r) return r;
    if (store) {
      cF = cF < NFOCK ? cF + 1 : NFOCK;
      counter = (counter + 1) % NFOCK;
      r = pulay_store(counter, cF);
      if (r) return r;
    }
    prof_mark("finish+host_sync", st);
    SEQM_CUDA(cudaMemsetAsync(wk.stat, 0, 16, st));
    finish_step_kernel<<<cdiv(B, 256), 256, 0, st>>>(B, opt->eps, wk.Enew, Eelec, wk.err, nc, niter, wk.stat);
    SEQM_CHECK_LAUNCH("finish_step_kernel");
    SEQM_CUDA(cudaMemcpyAsync(hstat, wk.stat, 16, cudaMemcpyDeviceToHost, st));
    SEQM_CUDA(cudaStreamSynchronize(st));
    left = (long long)hstat[0];
    if (opt->debug) {
      double mx;
      long long bits = (long long)hstat[1];
      memcpy(&mx, &bits, 8);
      printf("scf  %d %.17g %lld\n", k, mx, left);
      fflush(stdout);
    }
    ++k;
    return SEQM_OK;
  };
  auto mix = [&](int mode) -> int {
    prof_mark("mix", st);
    if (mode == MIX_ADAPTIVE) {
      mix_fac_kernel<<<B, 256, 0, st>>>(n4, nc, wk.Pnew, P, wk.Pold, wk.fac, nheavy, nhydro);
      SEQM_CHECK_LAUNCH("mix_fac_kernel");
    }
    mix_kernel<<<(unsigned)((size_t)B * chunks), 256, 0, st>>>(n4, chunks, mode, opt->alpha, nc, wk.Pnew, P, wk.Pold, wk.fac,
                                                              nheavy, nhydro);
    SEQM_CHECK_LAUNCH("mix_kernel");
    return SEQM_OK;
  };
#define SEQM_TRY(x) do { rc = (x); if (rc) return rc; } while (0)

  int counter = -1, cF = 0;
  if (opt->converger == 0) {
    while (left > 0) {
      SEQM_TRY(density(wk.Pnew));
      SEQM_TRY(mix(MIX_CONST));
      SEQM_TRY(refock_and_finish(false, counter, cF));
      if (k >= max_iter) break;
    }
  } else {
    bool done = false;
    for (int i = 0; i < 2 && !done; ++i) {
      if (left == 0) { done = true; break; }
      SEQM_TRY(density(wk.Pnew));
      SEQM_TRY(mix(MIX_DIRECT));
      SEQM_TRY(refock_and_finish(false, counter, cF));
    }
    if (!done && opt->converger == 1) {
      while (left > 0) {
        SEQM_TRY(density(wk.Pnew));
        SEQM_TRY(mix(MIX_ADAPTIVE));
        SEQM_TRY(refock_and_finish(false, counter, cF));
        if (k >= max_iter) break;
      }
    } else if (!done) {
      if (left > 0) {
        SEQM_TRY(density(wk.Pnew));
        SEQM_TRY(mix(MIX_ADAPTIVE));
        SEQM_TRY(refock_and_finish(false, counter, cF));
      }
      // two stores with direct steps (scf_loop.py:408-453): store uses the current F, P
      while (cF < 2 && left > 0) {
        cF = cF < NFOCK ? cF + 1 : NFOCK;
        counter = (counter + 1) % NFOCK;
        SEQM_TRY(pulay_store(counter, cF));
        SEQM_TRY(density(P));
        SEQM_TRY(refock_and_finish(false, counter, cF));
      }
      while (left > 0) {
        prof_mark("pulay_extrapolate", st);
        pulay_coeff_kernel<<<cdiv(B, 128), 128, 0, st>>>(B, counter, cF, nc, wk.EMAT, wk.coeff);
        if (packed_comm) {
          const int pchunks = red_chunks((size_t)mld * mld);
          pulay_extrapolate_packed_kernel<<<(unsigned)((size_t)B * pchunks), 256, 0, st>>>(n4, mld, pchunks, cF, nc, nheavy, nhydro,
                                                                                        wk.coeff, wk.FOCK, F);
        } else {
          pulay_extrapolate_kernel<<<(unsigned)((size_t)B * chunks), 256, 0, st>>>(n4, chunks, cF, nc, wk.coeff, wk.FOCK, F);
        }
        g_launches += 2;
        { cudaError_t e_ = cudaGetLastError(); if (e_ != cudaSuccess) { set_error(std::string("pulay_extrapolate_kernel: ") + cudaGetErrorString(e_)); return SEQM_ECUDA; } }
        SEQM_TRY(density(P));
        SEQM_TRY(refock_and_finish(true, counter, cF));
        if (k >= max_iter) break;
      }
    }
  }
  prof_mark("end", st);
  if (iterations_out) { iterations_out[0] = k; iterations_out[1] = (int32_t)left; }
  return SEQM_OK;
}

}  // extern "C"
