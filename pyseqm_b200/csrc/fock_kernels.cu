// K2: Fock build and electronic energy.
//   fock_pair_kernel<CLS>  one thread per pair of a class: exchange block F_AB = M_AB - 1/2 sum P_AB w,
//                          written to both triangles, and the two Coulomb vectors (w pB, pA w) of the pair
//                          to a per-pair scratch row (no atomics)
//   fock_diag_kernel       one warp per atom: one-centre terms + fixed-order sum of its partners' Coulomb
//                          vectors + core block, written to both triangles of the diagonal block
//   elec_energy_kernel     one CTA per molecule, 0.5 sum P.(H+F)   (or the XL-BOMD shadow energy)
// reference: seqm/seqm_functions/fock.py:6-139, energy.py:19-36, XLBOMD.py:40-52
#include <mutex>

#include "abi_common.cuh"
#include "dgemm.cuh"

namespace seqm {

#define KM(k) ((k) == 0 ? 0 : ((k) <= 2 ? 1 : ((k) <= 5 ? 2 : 3)))
#define KN(k) ((k) - KM(k) * (KM(k) + 1) / 2)
#define IND(a, b) ((a) >= (b) ? (a) * ((a) + 1) / 2 + (b) : (b) * ((b) + 1) / 2 + (a))
#define WT(k) ((KM(k) == KN(k)) ? 1.0 : 2.0)

template <int CLS>
__global__ void __launch_bounds__(128)
fock_pair_kernel(seqm_topology_t t, const double* __restrict__ wc, const double* __restrict__ P,
                 const double* __restrict__ M, const uint8_t* __restrict__ active, double* __restrict__ F,
                 double* __restrict__ pairvec) {
  const int slot = blockIdx.x * blockDim.x + threadIdx.x;
  if (slot >= t.ncls[CLS]) return;
  const int p = t.cls_pairs[CLS][slot];
  const int i = t.idxi[p], j = t.idxj[p];
  const int mol = t.atom_mol[i];
  if (active && !active[mol]) return;
  const int S = (int)t.molsize;
  const size_t n4 = (size_t)4 * S;
  const int si = t.atom_slot[i], sj = t.atom_slot[j];
  const double* Pm = P + (size_t)mol * n4 * n4;
  double* Fm = F + (size_t)mol * n4 * n4;
  const double* Mb = M + (((size_t)mol * S + si) * S + sj) * 16;
  double* pv = pairvec + (size_t)p * 20;
  constexpr int NA = (CLS == SEQM_CLS_HH) ? 1 : 10;
  constexpr int NB = (CLS == SEQM_CLS_XX) ? 10 : 1;
  constexpr int OA = (CLS == SEQM_CLS_HH) ? 1 : 4;
  constexpr int OB = (CLS == SEQM_CLS_XX) ? 4 : 1;
  double pA[NA], pB[NB];
  {
    const double* a = Pm + ((size_t)4 * si) * n4 + 4 * si;
    const double* b = Pm + ((size_t)4 * sj) * n4 + 4 * sj;
#pragma unroll
    for (int k = 0; k < NA; ++k) pA[k] = WT(k) * a[(size_t)KN(k) * n4 + KM(k)];
#pragma unroll
    for (int l = 0; l < NB; ++l) pB[l] = WT(l) * b[(size_t)KN(l) * n4 + KM(l)];
  }
  double Pab[OA][OB];
#pragma unroll
  for (int mu = 0; mu < OA; ++mu)
#pragma unroll
    for (int nu = 0; nu < OB; ++nu) Pab[mu][nu] = Pm[((size_t)4 * si + mu) * n4 + 4 * sj + nu];
  double toA[NA], toB[NB], Fab[OA][OB];
#pragma unroll
  for (int k = 0; k < NA; ++k) toA[k] = 0.0;
#pragma unroll
  for (int l = 0; l < NB; ++l) toB[l] = 0.0;
#pragma unroll
  for (int mu = 0; mu < OA; ++mu)
#pragma unroll
    for (int nu = 0; nu < OB; ++nu) Fab[mu][nu] = 0.0;
  const double* wp = (CLS == SEQM_CLS_HH) ? wc + slot : wc + ((size_t)(slot >> 5) * (NA * NB)) * 32 + (slot & 31);
#pragma unroll
  for (int k = 0; k < NA; ++k) {
#pragma unroll
    for (int l = 0; l < NB; ++l) {
      const double wv = (CLS == SEQM_CLS_HH) ? wp[0] : wp[(size_t)(k * NB + l) * 32];
      toA[k] += wv * pB[l];
      toB[l] += pA[k] * wv;
      // exchange: every ordered (mu,nu) with IND = k and (la,sg) with IND = l
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
      Fm[((size_t)4 * si + mu) * n4 + 4 * sj + nu] = f;
      Fm[((size_t)4 * sj + nu) * n4 + 4 * si + mu] = f;
    }
}

__global__ void __launch_bounds__(256)
fock_diag_kernel(seqm_topology_t t, const double* __restrict__ P, const double* __restrict__ M,
                 const double* __restrict__ gss, const double* __restrict__ gpp, const double* __restrict__ gsp,
                 const double* __restrict__ gp2, const double* __restrict__ hsp,
                 const uint8_t* __restrict__ active, const double* __restrict__ pairvec, double* __restrict__ F) {
  const int a = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (a >= t.natoms) return;
  const int mol = t.atom_mol[a];
  if (active && !active[mol]) return;
  const bool heavy = t.Z[a] > 1;
  double acc[10];
#pragma unroll
  for (int k = 0; k < 10; ++k) acc[k] = 0.0;
  for (int q = t.start_i[a] + lane; q < t.start_i[a + 1]; q += 32) {
    const double* pv = pairvec + (size_t)q * 20;
    acc[0] += pv[0];
    if (heavy) {
#pragma unroll
      for (int k = 1; k < 10; ++k) acc[k] += pv[k];
    }
  }
  for (int q = t.start_j[a] + lane; q < t.start_j[a + 1]; q += 32) {
    const double* pv = pairvec + (size_t)t.perm_j[q] * 20 + 10;
    acc[0] += pv[0];
    if (heavy) {
#pragma unroll
      for (int k = 1; k < 10; ++k) acc[k] += pv[k];
    }
  }
#pragma unroll
  for (int k = 0; k < 10; ++k) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc[k] += __shfl_xor_sync(0xffffffffu, acc[k], o);
  }
  if (lane == 0) {
    const int S = (int)t.molsize;
    const size_t n4 = (size_t)4 * S;
    const int s = t.atom_slot[a];
    const double* Pd = P + (size_t)mol * n4 * n4 + ((size_t)4 * s) * n4 + 4 * s;
    double* Fd = F + (size_t)mol * n4 * n4 + ((size_t)4 * s) * n4 + 4 * s;
    const double* Mb = M + (((size_t)mol * S + s) * S + s) * 16;
    const double GSS = gss[a], GPP = gpp[a], GSP = gsp[a], GP2 = gp2[a], HSP = hsp[a];
    double Pu[10];
#pragma unroll
    for (int k = 0; k < 10; ++k) Pu[k] = Pd[(size_t)KN(k) * n4 + KM(k)];
    const double Pss = Pu[0], Pxx = Pu[2], Pyy = Pu[5], Pzz = Pu[9];
    const double Ppp = Pxx + Pyy + Pzz;
    double one[10];
    one[0] = 0.5 * Pss * GSS + Ppp * (GSP - 0.5 * HSP);
#pragma unroll
    for (int k = 1; k < 10; ++k) {
      const int m = KM(k), n = KN(k);
      if (m == n) one[k] = Pss * (GSP - 0.5 * HSP) + 0.5 * Pu[k] * GPP + (Ppp - Pu[k]) * (1.25 * GP2 - 0.25 * GPP);
      else if (n == 0) one[k] = Pu[k] * (1.5 * HSP - 0.5 * GSP);
      else one[k] = Pu[k] * (0.75 * GPP - 1.25 * GP2);
    }
#pragma unroll
    for (int k = 0; k < 10; ++k) {
      const int m = KM(k), n = KN(k);
      const double f = Mb[n * 4 + m] + one[k] + acc[k];
      Fd[(size_t)n * n4 + m] = f;
      if (m != n) Fd[(size_t)m * n4 + n] = f;
    }
  }
}

// E[b] = 0.5 sum_ij P_ij (H_ij + F_ij)       (xl == 0)
// E[b] = sum_ij D_ij F_ij - 0.5 (F_ij - H_ij) P_ij   (xl != 0)
__global__ void __launch_bounds__(256)
elec_energy_kernel(int S, const double* __restrict__ P, const double* __restrict__ F, const double* __restrict__ M,
                   const double* __restrict__ D, int xl, const uint8_t* __restrict__ active, double* __restrict__ E) {
  const int mol = blockIdx.x;
  if (active && !active[mol]) return;
  const int n4 = 4 * S;
  const size_t base = (size_t)mol * n4 * n4;
  double acc = 0.0;
  for (int idx = threadIdx.x; idx < n4 * n4; idx += blockDim.x) {
    const int i = idx / n4, j = idx - i * n4;
    const int ai = i >> 2, mu = i & 3, aj = j >> 2, nu = j & 3;
    double h;
    if (ai < aj || (ai == aj && mu <= nu)) h = M[(((size_t)mol * S + ai) * S + aj) * 16 + mu * 4 + nu];
    else h = M[(((size_t)mol * S + aj) * S + ai) * 16 + nu * 4 + mu];
    const double f = F[base + idx], pv = P[base + idx];
    if (xl) acc += D[base + idx] * f - 0.5 * (f - h) * pv;
    else acc += 0.5 * pv * (h + f);
  }
  __shared__ double red[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    for (int q = 0; q < (int)(blockDim.x >> 5); ++q) s += red[q];
    E[mol] = s;
  }
}

// Large molecules (one CTA per molecule would leave the GPU idle: config 5 is ONE molecule with a 6000 x 6000 padded
// matrix): the same sum cut into chunks of RED_CHUNK elements, one CTA per (molecule, chunk), partial sums in
// caller-provided scratch, summed in a fixed order by elec_energy_sum_kernel (deterministic, so that replicated
// ranks of the tile-sharded SP2 take identical convergence decisions).
__global__ void __launch_bounds__(256)
elec_energy_part_kernel(int S, int chunks, const double* __restrict__ P, const double* __restrict__ F,
                        const double* __restrict__ M, const double* __restrict__ D, int xl,
                        const uint8_t* __restrict__ active, double* __restrict__ part,
                        const int32_t* __restrict__ nheavy, const int32_t* __restrict__ nhydro) {
  const int mol = blockIdx.x / chunks, chunk = blockIdx.x - mol * chunks;
  if (active && !active[mol]) return;
  const int n4 = 4 * S;
  // nheavy / nhydro given: only the rows that hold an orbital are read (the others are zero in P): for hydrogen-rich
  // molecules that is half of the padded matrix or less
  const int nh4 = nheavy ? 4 * nheavy[mol] : n4;
  const int mrow = nheavy ? nh4 + nhydro[mol] : n4;
  const size_t nn = (size_t)mrow * n4, base = (size_t)mol * n4 * n4;
  const size_t lo = (size_t)chunk * RED_CHUNK;
  const size_t hi = lo + RED_CHUNK < nn ? lo + RED_CHUNK : nn;
  double acc = 0.0;
  for (size_t idx0 = lo + threadIdx.x; idx0 < hi; idx0 += blockDim.x) {
    int ip, j;
    row_col(idx0, n4, ip, j);
    const int i = ip < nh4 ? ip : nh4 + 4 * (ip - nh4);
    const size_t idx = (size_t)i * n4 + j;
    const int ai = i >> 2, mu = i & 3, aj = j >> 2, nu = j & 3;
    double h;
    if (ai < aj || (ai == aj && mu <= nu)) h = M[(((size_t)mol * S + ai) * S + aj) * 16 + mu * 4 + nu];
    else h = M[(((size_t)mol * S + aj) * S + ai) * 16 + nu * 4 + mu];
    const double f = F[base + idx], pv = P[base + idx];
    if (xl) acc += D[base + idx] * f - 0.5 * (f - h) * pv;
    else acc += 0.5 * pv * (h + f);
  }
  __shared__ double red[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s_ = 0.0;
    for (int q = 0; q < (int)(blockDim.x >> 5); ++q) s_ += red[q];
    part[(size_t)mol * chunks + chunk] = s_;
  }
}

// out[b * ostride] = sum of the `chunks` partial sums of item b, fixed order; one CTA per item
__global__ void __launch_bounds__(256)
chunk_sum_kernel(int chunks, const uint8_t* __restrict__ active, int items_per_mol, const double* __restrict__ part,
                 double* __restrict__ out, const int32_t* __restrict__ out_index) {
  const int b = blockIdx.x;
  if (active && !active[b / items_per_mol]) return;
  double acc = 0.0;
  for (int q = threadIdx.x; q < chunks; q += blockDim.x) acc += part[(size_t)b * chunks + q];
  __shared__ double red[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s_ = 0.0;
    for (int q = 0; q < (int)(blockDim.x >> 5); ++q) s_ += red[q];
    out[out_index ? out_index[b] : b] = s_;
  }
}

int red_chunks(size_t nn) { return (int)((nn + RED_CHUNK - 1) / RED_CHUNK); }

int launch_chunk_sum(int items, int chunks, const uint8_t* active, int items_per_mol, const double* part, double* out,
                     const int32_t* out_index, cudaStream_t st) {
  chunk_sum_kernel<<<items, 256, 0, st>>>(chunks, active, items_per_mol, part, out, out_index);
  SEQM_CHECK_LAUNCH("chunk_sum_kernel");
  return SEQM_OK;
}

int launch_elec_energy(int64_t nmol, int64_t molsize, const double* P, const double* F, const double* M,
                       const double* D, int xl, const uint8_t* active, double* E, double* scratch, cudaStream_t st,
                       const int32_t* nheavy, const int32_t* nhydro) {
  const size_t nn = (size_t)16 * molsize * molsize;
  const int chunks = red_chunks(nn);
  if (!scratch || chunks == 1) {
    elec_energy_kernel<<<(unsigned)nmol, 256, 0, st>>>((int)molsize, P, F, M, D, xl, active, E);
    SEQM_CHECK_LAUNCH("elec_energy_kernel");
    return SEQM_OK;
  }
  SEQM_REQUIRE((long long)nmol * chunks < (1LL << 31), "elec_energy: too many chunks");
  elec_energy_part_kernel<<<(unsigned)(nmol * chunks), 256, 0, st>>>((int)molsize, chunks, P, F, M, D, xl, active, scratch,
                                                                   nheavy, nhydro);
  SEQM_CHECK_LAUNCH("elec_energy_part_kernel");
  return launch_chunk_sum((int)nmol, chunks, active, 1, scratch, E, nullptr, st);
}

int launch_fock(const seqm_topology_t& t, seqm_w_t w, const double* P, const double* M, const double* g_ss,
                const double* g_pp, const double* g_sp, const double* g_p2, const double* h_sp,
                const uint8_t* active, double* F, double* Eelec, double* pairvec, double* scratch, cudaStream_t st,
                const int32_t* nheavy, const int32_t* nhydro) {
  prof_mark("fock_pair", st);
  if (t.ncls[2] > 0) {
    fock_pair_kernel<SEQM_CLS_XX><<<cdiv(t.ncls[2], 128), 128, 0, st>>>(t, w.wXX, P, M, active, F, pairvec);
    SEQM_CHECK_LAUNCH("fock_pair_kernel<XX>");
  }
  if (t.ncls[1] > 0) {
    fock_pair_kernel<SEQM_CLS_XH><<<cdiv(t.ncls[1], 128), 128, 0, st>>>(t, w.wXH, P, M, active, F, pairvec);
    SEQM_CHECK_LAUNCH("fock_pair_kernel<XH>");
  }
  if (t.ncls[0] > 0) {
    fock_pair_kernel<SEQM_CLS_HH><<<cdiv(t.ncls[0], 128), 128, 0, st>>>(t, w.wHH, P, M, active, F, pairvec);
    SEQM_CHECK_LAUNCH("fock_pair_kernel<HH>");
  }
  prof_mark("fock_diag", st);
  if (t.natoms > 0) {
    fock_diag_kernel<<<cdiv(t.natoms * 32, 256), 256, 0, st>>>(t, P, M, g_ss, g_pp, g_sp, g_p2, h_sp, active, pairvec, F);
    SEQM_CHECK_LAUNCH("fock_diag_kernel");
  }
  prof_mark("elec_energy", st);
  if (Eelec && t.nmol > 0) {
    const int rc_ = launch_elec_energy(t.nmol, t.molsize, P, F, M, nullptr, 0, active, Eelec, scratch, st, nheavy, nhydro);
    if (rc_) return rc_;
  }
  prof_mark("end", st);
  return SEQM_OK;
}

// ---- allocation-free variant of the chunked energy sum (seqm_elec_energy_noalloc) -----------------------------------
// Its partial sums live in a file-scope device array instead of caller scratch: 512 KiB that every GPU loading the
// library holds until exit.  All calls on a device share the array: calls that overlap on the GPU (different streams)
// race on it.  Nothing in it has to start at zero - every partial sum a call reads was written by that same call.
constexpr size_t NOALLOC_MAX_PARTIALS = SEQM_NOALLOC_MAX_PARTIALS;
static __device__ __align__(256) double g_noalloc_part[NOALLOC_MAX_PARTIALS];

// Device address of g_noalloc_part on the CURRENT device.  In host code the symbol's name is only a host-side stand-in,
// not a device address, so it is resolved with cudaGetSymbolAddress - once per device, with that device current.
static int noalloc_scratch(double** out) {
  constexpr int MAXDEV = 64;
  static std::mutex mtx;
  static double* ptrs[MAXDEV] = {nullptr};
  int dev = -1;
  SEQM_CUDA(cudaGetDevice(&dev));
  SEQM_REQUIRE(dev >= 0 && dev < MAXDEV, "seqm_elec_energy_noalloc: device ordinal out of range");
  std::lock_guard<std::mutex> lock(mtx);
  if (!ptrs[dev]) {
    double* p = nullptr;
    SEQM_CUDA(cudaGetSymbolAddress((void**)&p, g_noalloc_part));
    SEQM_REQUIRE(p != nullptr, "seqm_elec_energy_noalloc: cudaGetSymbolAddress returned a null address");
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return SEQM_OK;
}

}  // namespace seqm

using namespace seqm;

extern "C" {

int seqm_fock(const seqm_topology_t* topo, seqm_w_t w, const double* P, const double* M, const double* g_ss,
              const double* g_pp, const double* g_sp, const double* g_p2, const double* h_sp,
              const uint8_t* active, double* F, double* Eelec, double* pairvec, void* work, void* stream) {
  SEQM_REQUIRE(topo && P && M && g_ss && g_pp && g_sp && g_p2 && h_sp && F && pairvec, "seqm_fock: null pointer");
  return launch_fock(*topo, w, P, M, g_ss, g_pp, g_sp, g_p2, h_sp, active, F, Eelec, pairvec, (double*)work,
                     (cudaStream_t)stream);
}

int seqm_elec_energy(int64_t nmol, int64_t molsize, const double* P, const double* F, const double* M,
                     const double* D, int xl, const uint8_t* active, double* E, void* work, void* stream) {
  SEQM_REQUIRE(P && F && M && E && nmol >= 0 && molsize > 0 && molsize < 16384, "seqm_elec_energy: bad argument");
  SEQM_REQUIRE(!xl || D, "seqm_elec_energy: xl needs D");
  if (nmol == 0) return SEQM_OK;
  return launch_elec_energy(nmol, molsize, P, F, M, D, xl, active, E, (double*)work, (cudaStream_t)stream);
}

int seqm_elec_energy_noalloc(int64_t nmol, int64_t molsize, const double* P, const double* F, const double* M,
                             const double* D, int xl, const uint8_t* active, double* E, void* stream) {
  SEQM_REQUIRE(P && F && M && E && nmol >= 0 && molsize > 0 && molsize < 16384, "seqm_elec_energy_noalloc: bad argument");
  SEQM_REQUIRE(!xl || D, "seqm_elec_energy_noalloc: xl needs D");
  if (nmol == 0) return SEQM_OK;
  const size_t nn = (size_t)16 * molsize * molsize;
  const int chunks = red_chunks(nn);
  if ((unsigned long long)nmol * (unsigned long long)chunks > (unsigned long long)NOALLOC_MAX_PARTIALS) {
    char msg[256];
    snprintf(msg, sizeof msg,
             "seqm_elec_energy_noalloc: %lld molecules x %d chunks of 8192 matrix elements = %lld partial sums exceed the "
             "static scratch of %lld (SEQM_NOALLOC_MAX_PARTIALS); use seqm_elec_energy with caller scratch",
             (long long)nmol, chunks, (long long)nmol * chunks, (long long)NOALLOC_MAX_PARTIALS);
    return fail(SEQM_EINVAL, msg);
  }
  double* part = nullptr;
  const int rc = noalloc_scratch(&part);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  prof_mark("elec_energy", st);
  // always the chunked pair of kernels (also when a molecule is a single chunk): partial sums -> fixed-order sum
  elec_energy_part_kernel<<<(unsigned)(nmol * chunks), 256, 0, st>>>((int)molsize, chunks, P, F, M, D, xl, active, part, nullptr, nullptr);
  SEQM_CHECK_LAUNCH("elec_energy_part_kernel");
  const int rc2 = launch_chunk_sum((int)nmol, chunks, active, 1, part, E, nullptr, st);
  prof_mark("end", st);
  return rc2;
}

size_t seqm_elec_energy_workspace(int64_t nmol, int64_t molsize) {
  return (size_t)nmol * red_chunks((size_t)16 * molsize * molsize) * 8;
}

}  // extern "C"
