// K5: the elementwise half of an XL-BOMD / velocity-Verlet step as a few fused kernels, so that one MD step is
// nothing but launches of this library (and can be captured in ONE CUDA graph):
//   xl_extrapolate_kernel   P = kappa D + sum_j c[(cidx + j)] Pt[j],  Pt[m-1-cidx] = P        (XLBOMD.py:293-295)
//   verlet_kick_drift_kernel  v += 0.5 a dt ; x += v dt                                         (XLBOMD.py:285-286)
//   verlet_finish_kernel    force = -gR ; a = force / mass * acc_scale ; v += 0.5 a dt ; step counter += 1 (:302-305)
//   energy_finish_kernel    Enuc[b] = sum of the molecule's pair core repulsions (fixed order), Etot = Eelec + Enuc,
//                           Hf = Etot + hconst[b]   (energy.py:92-94,109-118 with the per-molecule constants folded)
//   md_thermo_kernel        kinetic energy, temperature, atomic charges q = Z' - sum_mu P_mumu and dipole per molecule
//                           (MolecularDynamics.py:229-243)
// The arithmetic is written with explicit round-to-nearest multiplies and adds in the ORDER of the reference's tensor
// expressions (no FMA contraction), so a trajectory is bit-identical to the eager-torch form of the same step.
// The step index lives on the device (`step_dev`), which is what lets one captured graph serve every step.
// reference: seqm/XLBOMD.py:277-307, seqm/MolecularDynamics.py:229-273
#include "abi_common.cuh"

namespace seqm {

constexpr int XL_MAXM = 10;
struct XlCoeff { double c[2 * XL_MAXM]; };

__global__ void __launch_bounds__(256)
xl_extrapolate_kernel(long long nelem, int mm, double kappa, XlCoeff cf, const long long* __restrict__ step_dev,
                      long long step_host, const double* __restrict__ D, double* __restrict__ Pt,
                      double* __restrict__ P) {
  const long long step = step_dev ? *step_dev : step_host;
  const int cidx = (int)(step % mm);
  double* slot = Pt + (size_t)(mm - 1 - cidx) * nelem;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < nelem;
       idx += (long long)gridDim.x * blockDim.x) {
    // torch: coeff_D * D + torch.sum(coeff[cidx:cidx+m].reshape(-1,1,1,1) * Pt, dim=0): the sum over j runs in order
    double s = __dmul_rn(cf.c[cidx], Pt[idx]);
    for (int j = 1; j < mm; ++j) s = __dadd_rn(s, __dmul_rn(cf.c[cidx + j], Pt[(size_t)j * nelem + idx]));
    const double p = __dadd_rn(__dmul_rn(kappa, D[idx]), s);
    P[idx] = p;
    slot[idx] = p;
  }
}

__global__ void verlet_kick_drift_kernel(long long n, double dt, const double* __restrict__ acc,
                                         double* __restrict__ v, double* __restrict__ x) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double vn = __dadd_rn(v[i], __dmul_rn(__dmul_rn(0.5, acc[i]), dt));      // velocities.add_(0.5 * acc * dt)
  v[i] = vn;
  x[i] = __dadd_rn(x[i], __dmul_rn(vn, dt));                                      // coordinates.add_(velocities * dt)
}

__global__ void verlet_finish_kernel(long long n, double dt, double acc_scale, const double* __restrict__ gR,
                                     const double* __restrict__ mass, double* __restrict__ force,
                                     double* __restrict__ acc, double* __restrict__ v, long long* __restrict__ step_dev) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i == 0 && step_dev) *step_dev += 1;
  if (i >= n) return;
  const double f = -gR[i];
  const double a = __dmul_rn(__ddiv_rn(f, mass[i / 3]), acc_scale);               // force / mass * acc_scale
  force[i] = f;
  acc[i] = a;
  v[i] = __dadd_rn(v[i], __dmul_rn(__dmul_rn(0.5, a), dt));
}

// one warp per molecule; pair_off[b] .. pair_off[b+1] = the molecule's pairs (the Parser orders pairs by molecule)
__global__ void __launch_bounds__(256)
energy_finish_kernel(int nmol, const int32_t* __restrict__ pair_off, const double* __restrict__ EnucAB,
                     const double* __restrict__ Eelec, const double* __restrict__ hconst, double* __restrict__ Enuc,
                     double* __restrict__ Etot, double* __restrict__ Hf) {
  const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (b >= nmol) return;
  double s = 0.0;
  for (int p = pair_off[b] + lane; p < pair_off[b + 1]; p += 32) s += EnucAB[p];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) {
    const double et = Eelec[b] + s;
    Enuc[b] = s;
    Etot[b] = et;
    Hf[b] = et + hconst[b];
  }
}

// one CTA per molecule: Ek = sum 0.5 m v^2 * scale, T, q[a] = tore[a] - sum_mu P[4a+mu][4a+mu], dipole = sum q x
__global__ void __launch_bounds__(128)
md_thermo_kernel(int S, double ke_scale, double ev_to_kelvin, const double* __restrict__ mass,
                 const double* __restrict__ tore, const int32_t* __restrict__ natoms, const double* __restrict__ v,
                 const double* __restrict__ x, const double* __restrict__ P, double* __restrict__ Ek,
                 double* __restrict__ T, double* __restrict__ q, double* __restrict__ dip) {
  const int b = blockIdx.x, tid = threadIdx.x, n4 = 4 * S;
  __shared__ double red[4][4];
  double ek = 0.0, d0 = 0.0, d1 = 0.0, d2 = 0.0;
  for (int a = tid; a < S; a += blockDim.x) {
    const size_t ia = (size_t)b * S + a;
    const double m = mass[ia];
    const double* va = v + ia * 3;
    ek += 0.5 * m * va[0] * va[0] + 0.5 * m * va[1] * va[1] + 0.5 * m * va[2] * va[2];
    if (P) {
      const double* Pb = P + (size_t)b * n4 * n4;
      double pop = 0.0;
      for (int mu = 0; mu < 4; ++mu) pop += Pb[(size_t)(4 * a + mu) * n4 + 4 * a + mu];
      const double qa = tore[ia] - pop;
      q[ia] = qa;
      d0 += qa * x[ia * 3];
      d1 += qa * x[ia * 3 + 1];
      d2 += qa * x[ia * 3 + 2];
    }
  }
  double vals[4] = {ek, d0, d1, d2};
#pragma unroll
  for (int k = 0; k < 4; ++k) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) vals[k] += __shfl_xor_sync(0xffffffffu, vals[k], o);
    if ((tid & 31) == 0) red[k][tid >> 5] = vals[k];
  }
  __syncthreads();
  if (tid == 0) {
    double out[4];
    for (int k = 0; k < 4; ++k) {
      out[k] = 0.0;
      for (int w = 0; w < (int)(blockDim.x >> 5); ++w) out[k] += red[k][w];
    }
    const double e = out[0] * ke_scale;
    Ek[b] = e;
    T[b] = e * ev_to_kelvin / (1.5 * (double)natoms[b]);
    if (P) { dip[b * 3] = out[1]; dip[b * 3 + 1] = out[2]; dip[b * 3 + 2] = out[3]; }
  }
}

}  // namespace seqm

using namespace seqm;

extern "C" {

int seqm_xl_extrapolate(int64_t nelem, int m, double kappa, const double* coeff2m, const int64_t* step_dev,
                        int64_t step_host, const double* D, double* Pt, double* P, void* stream) {
  SEQM_REQUIRE(nelem >= 0 && m >= 1 && m <= XL_MAXM && coeff2m && D && Pt && P, "seqm_xl_extrapolate: bad argument");
  if (nelem == 0) return SEQM_OK;
  XlCoeff cf;
  for (int i = 0; i < 2 * m; ++i) cf.c[i] = coeff2m[i];      // HOST array (the 2m doubled coefficients of XLBOMD.py:259-262)
  cudaStream_t st = (cudaStream_t)stream;
  prof_mark("xl_extrapolate", st);
  long long blocks = (nelem + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  xl_extrapolate_kernel<<<(unsigned)blocks, 256, 0, st>>>((long long)nelem, m, kappa, cf, (const long long*)step_dev,
                                                         (long long)step_host, D, Pt, P);
  SEQM_CHECK_LAUNCH("xl_extrapolate_kernel");
  prof_mark("end", st);
  return SEQM_OK;
}

int seqm_verlet_kick_drift(int64_t n, double dt, const double* acc, double* v, double* x, void* stream) {
  SEQM_REQUIRE(n >= 0 && acc && v && x, "seqm_verlet_kick_drift: bad argument");
  if (n == 0) return SEQM_OK;
  cudaStream_t st = (cudaStream_t)stream;
  prof_mark("verlet", st);
  verlet_kick_drift_kernel<<<cdiv(n, 256), 256, 0, st>>>((long long)n, dt, acc, v, x);
  SEQM_CHECK_LAUNCH("verlet_kick_drift_kernel");
  prof_mark("end", st);
  return SEQM_OK;
}

int seqm_verlet_finish(int64_t n, double dt, double acc_scale, const double* gR, const double* mass, double* force,
                       double* acc, double* v, int64_t* step_dev, void* stream) {
  SEQM_REQUIRE(n >= 0 && gR && mass && force && acc && v, "seqm_verlet_finish: bad argument");
  if (n == 0) return SEQM_OK;
  cudaStream_t st = (cudaStream_t)stream;
  prof_mark("verlet", st);
  verlet_finish_kernel<<<cdiv(n, 256), 256, 0, st>>>((long long)n, dt, acc_scale, gR, mass, force, acc, v,
                                                    (long long*)step_dev);
  SEQM_CHECK_LAUNCH("verlet_finish_kernel");
  prof_mark("end", st);
  return SEQM_OK;
}

int seqm_energy_finish(int64_t nmol, const int32_t* pair_off, const double* EnucAB, const double* Eelec,
                       const double* hconst, double* Enuc, double* Etot, double* Hf, void* stream) {
  SEQM_REQUIRE(nmol >= 0 && pair_off && EnucAB && Eelec && hconst && Enuc && Etot && Hf, "seqm_energy_finish: bad argument");
  if (nmol == 0) return SEQM_OK;
  cudaStream_t st = (cudaStream_t)stream;
  prof_mark("energy_finish", st);
  energy_finish_kernel<<<cdiv(nmol, 8), 256, 0, st>>>((int)nmol, pair_off, EnucAB, Eelec, hconst, Enuc, Etot, Hf);
  SEQM_CHECK_LAUNCH("energy_finish_kernel");
  prof_mark("end", st);
  return SEQM_OK;
}

int seqm_md_thermo(int64_t nmol, int64_t molsize, double ke_scale, double ev_to_kelvin, const double* mass,
                   const double* tore, const int32_t* natoms, const double* v, const double* x, const double* P,
                   double* Ek, double* T, double* q, double* dipole, void* stream) {
  SEQM_REQUIRE(nmol >= 0 && molsize > 0 && mass && natoms && v && Ek && T, "seqm_md_thermo: bad argument");
  SEQM_REQUIRE(!P || (tore && x && q && dipole), "seqm_md_thermo: charges need tore, x, q, dipole");
  if (nmol == 0) return SEQM_OK;
  cudaStream_t st = (cudaStream_t)stream;
  prof_mark("md_thermo", st);
  md_thermo_kernel<<<(unsigned)nmol, 128, 0, st>>>((int)molsize, ke_scale, ev_to_kelvin, mass, tore, natoms, v, x, P, Ek, T,
                                                  q, dipole);
  SEQM_CHECK_LAUNCH("md_thermo_kernel");
  prof_mark("end", st);
  return SEQM_OK;
}

}  // extern "C"
