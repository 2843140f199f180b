// Kp: the Parser's pair list (reference seqm/basics.py:43-118) generated on the device, WITH the outer cutoff:
// no O(B S^2) boolean masks, int64 index temporaries or nonzero() compactions, and the class lists / CSR tables of
// seqm_topology_t come out of the same pass instead of a dozen torch index kernels.
//   count   thread per atom slot (mol, a): pairs (a, b > a) with both atoms real and |R_b - R_a| < cutoff, split by pair
//           class (HH / XH / XX), and the pairs (a' < a, a) in which the atom is the second one
//   (exclusive scans of the five count arrays: done by the caller, any scan will do)
//   fill    thread per atom slot: walks b again in the same order and writes, at the scanned offsets, the reference's
//           tensors (int64: mask, pair_molid, ni, nj, idxi, idxj; fp64: xij, rij) and the topology (int32 idxi, idxj,
//           pair class, slot inside the class list, the class lists) plus the pair id of (a, b) in a B x S x S table
//   second  thread per atom slot b: collects the ids of its pairs (a < b, b) in ascending a from that table: perm_j
// Pair order = (molecule, a, b) lexicographic = the order of the reference's boolean-mask selection (basics.py:96-109).
// The distance arithmetic is the reference's, operation by operation and without FMA contraction:
//   d = R_b - R_a ; d2 = (dx dx + dy dy) + dz dz ; d2 == 0 -> 1e-4 ; dist = sqrt(d2) ; keep = dist < cutoff ;
//   rij = dist * lcf ; xij = d / dist                                                     (basics.py:91-108)
#include "abi_common.cuh"

namespace seqm {

struct ParserArgs {
  int nmol, S;
  const int64_t* species;      // [B][S]
  const double* R;             // [B][S][3]
  double cutoff, lcf;
};

__device__ __forceinline__ bool pair_dist(const double* __restrict__ ra, const double* __restrict__ rb, double cutoff,
                                          double& dx, double& dy, double& dz, double& dist) {
  dx = __dsub_rn(rb[0], ra[0]);
  dy = __dsub_rn(rb[1], ra[1]);
  dz = __dsub_rn(rb[2], ra[2]);
  double d2 = __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
  if (d2 == 0.0) d2 = 1.0e-4;
  dist = __dsqrt_rn(d2);
  return dist < cutoff;
}

// cnt[0..2][t]: pairs of class HH / XH / XX whose FIRST atom is slot t; cnt[3][t]: pairs whose SECOND atom is t;
// cnt[4][t]: 1 if the slot holds a real atom
__global__ void __launch_bounds__(128)
parser_count_kernel(ParserArgs A, int32_t* __restrict__ cnt) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long n = (long long)A.nmol * A.S;
  if (t >= n) return;
  const int mol = (int)(t / A.S), a = (int)(t - (long long)mol * A.S);
  const int64_t* sp = A.species + (size_t)mol * A.S;
  const double* Rm = A.R + (size_t)mol * A.S * 3;
  const int za = (int)sp[a];
  int c[3] = {0, 0, 0}, second = 0;
  if (za > 0) {
    const int ha = za > 1 ? 1 : 0;
    double dx, dy, dz, dist;
    for (int b = a + 1; b < A.S; ++b) {
      const int zb = (int)sp[b];
      if (zb > 0 && pair_dist(Rm + 3 * a, Rm + 3 * b, A.cutoff, dx, dy, dz, dist)) ++c[ha + (zb > 1 ? 1 : 0)];
    }
    for (int b = 0; b < a; ++b) {
      if (sp[b] > 0 && pair_dist(Rm + 3 * b, Rm + 3 * a, A.cutoff, dx, dy, dz, dist)) ++second;
    }
  }
  cnt[t] = c[0];
  cnt[n + t] = c[1];
  cnt[2 * n + t] = c[2];
  cnt[3 * n + t] = second;
  cnt[4 * n + t] = za > 0 ? 1 : 0;
}

struct ParserOut {
  // exclusive scans over the atom slots (int64): first-atom pairs (all classes), per class, real atoms
  const int64_t *off_pair, *off_cls[3], *off_real;
  int64_t *mask, *pair_molid, *ni, *nj, *idxi64, *idxj64;      // the reference's int64 tensors [P]
  double *xij, *rij;
  int32_t *idxi, *idxj, *pair_cls, *pair_slot, *cls_pairs[3];
  int32_t* pairid;                                             // [B][S][S] scratch: id of pair (a, b) or -1
};

__global__ void __launch_bounds__(128)
parser_fill_kernel(ParserArgs A, ParserOut O) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long n = (long long)A.nmol * A.S;
  if (t >= n) return;
  const int mol = (int)(t / A.S), a = (int)(t - (long long)mol * A.S);
  const int64_t* sp = A.species + (size_t)mol * A.S;
  const double* Rm = A.R + (size_t)mol * A.S * 3;
  const int za = (int)sp[a];
  int32_t* prow = O.pairid + (size_t)t * A.S;
  if (za <= 0) {
    for (int b = a + 1; b < A.S; ++b) prow[b] = -1;
    return;
  }
  const int ha = za > 1 ? 1 : 0;
  long long p = O.off_pair[t];
  long long cur[3] = {O.off_cls[0][t], O.off_cls[1][t], O.off_cls[2][t]};
  const long long ia = O.off_real[t];
  double dx, dy, dz, dist;
  for (int b = a + 1; b < A.S; ++b) {
    const int zb = (int)sp[b];
    if (!(zb > 0 && pair_dist(Rm + 3 * a, Rm + 3 * b, A.cutoff, dx, dy, dz, dist))) { prow[b] = -1; continue; }
    const long long ib = O.off_real[(long long)mol * A.S + b];
    const int cls = ha + (zb > 1 ? 1 : 0);
    O.mask[p] = t * A.S + b;
    O.pair_molid[p] = mol;
    O.ni[p] = za;
    O.nj[p] = zb;
    O.idxi64[p] = ia;
    O.idxj64[p] = ib;
    O.xij[3 * p] = __ddiv_rn(dx, dist);
    O.xij[3 * p + 1] = __ddiv_rn(dy, dist);
    O.xij[3 * p + 2] = __ddiv_rn(dz, dist);
    O.rij[p] = __dmul_rn(dist, A.lcf);
    O.idxi[p] = (int32_t)ia;
    O.idxj[p] = (int32_t)ib;
    O.pair_cls[p] = cls;
    O.pair_slot[p] = (int32_t)cur[cls];
    O.cls_pairs[cls][cur[cls]] = (int32_t)p;
    ++cur[cls];
    prow[b] = (int32_t)p;
    ++p;
  }
}

// perm_j: for every atom slot b, the ids of the pairs (a < b, b) in ascending a, at the scanned offset of the slot
__global__ void __launch_bounds__(128)
parser_second_kernel(ParserArgs A, const int64_t* __restrict__ off_second, const int32_t* __restrict__ pairid,
                     int32_t* __restrict__ perm_j) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long n = (long long)A.nmol * A.S;
  if (t >= n) return;
  const int mol = (int)(t / A.S), b = (int)(t - (long long)mol * A.S);
  if (A.species[t] <= 0) return;
  long long q = off_second[t];
  const int32_t* pm = pairid + (size_t)mol * A.S * A.S;
  for (int a = 0; a < b; ++a) {
    const int32_t id = pm[(size_t)a * A.S + b];
    if (id >= 0) perm_j[q++] = id;
  }
}

}  // namespace seqm

using namespace seqm;

extern "C" {

int seqm_parser_count(int64_t nmol, int64_t molsize, const int64_t* species, const double* R, double cutoff,
                      double lcf, int32_t* counts5, void* stream) {
  SEQM_REQUIRE(nmol >= 0 && molsize > 0 && species && R && counts5, "seqm_parser_count: bad argument");
  if (nmol == 0) return SEQM_OK;
  cudaStream_t st = (cudaStream_t)stream;
  ParserArgs A{(int)nmol, (int)molsize, species, R, cutoff, lcf};
  prof_mark("parser", st);
  parser_count_kernel<<<cdiv(nmol * molsize, 128), 128, 0, st>>>(A, counts5);
  SEQM_CHECK_LAUNCH("parser_count_kernel");
  prof_mark("end", st);
  return SEQM_OK;
}

int seqm_parser_fill(int64_t nmol, int64_t molsize, const int64_t* species, const double* R, double cutoff, double lcf,
                     const int64_t* off_pair, const int64_t* off_hh, const int64_t* off_xh, const int64_t* off_xx,
                     const int64_t* off_second, const int64_t* off_real, int64_t* mask, int64_t* pair_molid,
                     int64_t* ni, int64_t* nj, int64_t* idxi64, int64_t* idxj64, double* xij, double* rij,
                     int32_t* idxi, int32_t* idxj, int32_t* pair_cls, int32_t* pair_slot, int32_t* cls_hh,
                     int32_t* cls_xh, int32_t* cls_xx, int32_t* perm_j, int32_t* pairid_scratch, void* stream) {
  SEQM_REQUIRE(nmol >= 0 && molsize > 0 && species && R && off_pair && off_hh && off_xh && off_xx && off_second &&
                   off_real && pairid_scratch,
               "seqm_parser_fill: bad argument");
  if (nmol == 0) return SEQM_OK;
  cudaStream_t st = (cudaStream_t)stream;
  ParserArgs A{(int)nmol, (int)molsize, species, R, cutoff, lcf};
  ParserOut O{};
  O.off_pair = off_pair; O.off_cls[0] = off_hh; O.off_cls[1] = off_xh; O.off_cls[2] = off_xx; O.off_real = off_real;
  O.mask = mask; O.pair_molid = pair_molid; O.ni = ni; O.nj = nj; O.idxi64 = idxi64; O.idxj64 = idxj64;
  O.xij = xij; O.rij = rij; O.idxi = idxi; O.idxj = idxj; O.pair_cls = pair_cls; O.pair_slot = pair_slot;
  O.cls_pairs[0] = cls_hh; O.cls_pairs[1] = cls_xh; O.cls_pairs[2] = cls_xx; O.pairid = pairid_scratch;
  prof_mark("parser", st);
  parser_fill_kernel<<<cdiv(nmol * molsize, 128), 128, 0, st>>>(A, O);
  SEQM_CHECK_LAUNCH("parser_fill_kernel");
  parser_second_kernel<<<cdiv(nmol * molsize, 128), 128, 0, st>>>(A, off_second, pairid_scratch, perm_j);
  SEQM_CHECK_LAUNCH("parser_second_kernel");
  prof_mark("end", st);
  return SEQM_OK;
}

}  // extern "C"
