// Per-atom-pair mathematics of the NDDO core Hamiltonian, templated on the scalar type
// (double for the forward kernels, Dual<4> for the force adjoint):
//   * local-frame two-centre two-electron integrals (generated, repp_gen.cuh)
//   * the molecular-frame rotation, written as a block tensor transform w = T L T^T with
//     T = diag(1, R, R(x)R) over the packed index groups {ss}, {p s}, {p p}
//     (reference: seqm/seqm_functions/two_elec_two_center_int.py:56-878 writes the products out)
//   * Slater-orbital overlaps of s/p shells with n = 1, 2 (reference: diat_overlap.py:3-393)
//   * core-core repulsion (reference: energy.py:38-78)
// Everything here is __host__ __device__ so tests can run the same source on the CPU
// (tests/hostcheck) — the product only ever calls it from CUDA kernels.
#pragma once
#include "common.cuh"

namespace seqm {

#include "repp_gen.cuh"

// per-atom record consumed by the pair kernels (built once per call by atom_pack_kernel)
struct AtomPar {
  double zs, zp, bs, bp;          // Slater exponents, resonance parameters
  AtomMultipole mp;               // dd, qq, rho0, rho1, rho2
  double alpha, tore;             // core repulsion exponent, valence charge
  double zs15, zs25, zp25;        // zs^1.5, zs^2.5, zp^2.5
  double gK[4], gL[4], gM[4];     // AM1/PM3 core-repulsion Gaussians (unused ones have K = 0)
  double pad_[2];
};
static_assert(sizeof(AtomPar) == 28 * 8, "AtomPar layout");

// packed-10 index groups: (p s) products and (p p) products
#define SEQM_PS(m) ((m) == 0 ? 1 : ((m) == 1 ? 3 : 6))
#define SEQM_PP(q) ((q) == 0 ? 2 : ((q) == 1 ? 4 : ((q) == 2 ? 5 : ((q) == 3 ? 7 : ((q) == 4 ? 8 : 9)))))
#define SEQM_QM(q) ((q) == 0 ? 0 : ((q) <= 2 ? 1 : 2))                       // first index of pair q
#define SEQM_QN(q) ((q) == 0 ? 0 : ((q) == 1 ? 0 : ((q) == 2 ? 1 : ((q) - 3))))  // second index

// Local axes of MOPAC's rotate: sigma axis X = -xij (from atom j to atom i), then Y, Z.
// R[mu][a]: component mu (x,y,z) of axis a (0 sigma, 1 pi, 2 pi').
// Branches follow two_elec_two_center_int.py:236-265 (z2 > 1e-5 test, sign(x1), |a x1|).
template <class S>
SEQM_HD void local_frame(const S* xij, S R[3][3]) {
  const S X0 = -xij[0], X1 = -xij[1], X2 = -xij[2];
  S z3(0.0);
  if (fabs(val(X2)) < 1.0) z3 = sqrt_(1.0 - X2 * X2);
  R[0][0] = X0; R[1][0] = X1; R[2][0] = X2;
  if (val(z3) > 1.0e-5) {
    const S a = 1.0 / z3;
    R[0][2] = -(a * X0 * X2);
    R[1][2] = -(a * X1 * X2);
    R[0][1] = (val(X0) >= 0.0) ? -(a * X1) : (a * X1);
    R[1][1] = abs_(a * X0);
  } else {
    R[0][2] = S(1.0);
    R[1][2] = S(0.0);
    R[0][1] = S(0.0);
    R[1][1] = S(1.0);
  }
  R[2][2] = z3;
  R[2][1] = S(0.0);
}

// T_q[(m,n)][(a,b)] = R[m][a] R[n][b] + (a != b) R[m][b] R[n][a]
template <class S>
SEQM_HD void build_tq(const S R[3][3], S Tq[6][6]) {
#pragma unroll
  for (int q = 0; q < 6; ++q) {
    const int m = SEQM_QM(q), n = SEQM_QN(q);
#pragma unroll
    for (int c = 0; c < 6; ++c) {
      const int a = SEQM_QM(c), b = SEQM_QN(c);
      S t = R[m][a] * R[n][b];
      if (a != b) t = t + R[m][b] * R[n][a];
      Tq[q][c] = t;
    }
  }
}

// heavy-heavy pair: all 100 molecular-frame integrals, delivered to sink(k, l, value)
template <class S, class Sink>
SEQM_HD void rotate_xx(const S* ri, const S R[3][3], Sink& sink) {
  S Tq[6][6];
  build_tq(R, Tq);
  sink(0, 0, ri[0]);
#pragma unroll
  for (int m = 0; m < 3; ++m) {
    sink(SEQM_PS(m), 0, R[m][0] * ri[1]);
    sink(0, SEQM_PS(m), R[m][0] * ri[4]);
#pragma unroll
    for (int n = 0; n < 3; ++n)
      sink(SEQM_PS(m), SEQM_PS(n), ri[5] * (R[m][0] * R[n][0]) + ri[6] * (R[m][1] * R[n][1] + R[m][2] * R[n][2]));
  }
#pragma unroll
  for (int q = 0; q < 6; ++q) {
    const S a0 = Tq[q][0], a1 = Tq[q][1], a2 = Tq[q][2], a3 = Tq[q][3], a4 = Tq[q][4], a5 = Tq[q][5];
    const S as = a2 + a5;
    sink(SEQM_PP(q), 0, a0 * ri[2] + as * ri[3]);
    sink(0, SEQM_PP(q), a0 * ri[10] + as * ri[11]);
    const S u0 = a0 * ri[7] + as * ri[8], u1 = a1 * ri[9], u2 = a3 * ri[9];
    const S v0 = a0 * ri[12] + as * ri[13], v1 = a1 * ri[14], v2 = a3 * ri[14];
#pragma unroll
    for (int l = 0; l < 3; ++l) {
      sink(SEQM_PP(q), SEQM_PS(l), u0 * R[l][0] + u1 * R[l][1] + u2 * R[l][2]);
      sink(SEQM_PS(l), SEQM_PP(q), v0 * R[l][0] + v1 * R[l][1] + v2 * R[l][2]);
    }
    // row q of Tq L_qq, then contract with every Tq row
    const S y0 = a0 * ri[15] + as * ri[16];
    const S y1 = a1 * ri[19];
    const S y2 = a0 * ri[17] + a2 * ri[18] + a5 * ri[20];
    const S y3 = a3 * ri[19];
    const S y4 = a4 * ri[21];
    const S y5 = a0 * ri[17] + a5 * ri[18] + a2 * ri[20];
#pragma unroll
    for (int p = 0; p < 6; ++p)
      sink(SEQM_PP(q), SEQM_PP(p),
           y0 * Tq[p][0] + y1 * Tq[p][1] + y2 * Tq[p][2] + y3 * Tq[p][3] + y4 * Tq[p][4] + y5 * Tq[p][5]);
  }
}

// ---------------------------------------------------------------------------------------------
// Reverse-mode adjoint of rotate_xx for the force kernel.  With w = T L T^T, T = diag(1, R, Tq(R)) and an upstream
// G = dE/dw, the energy E = sum_kl G_kl w_kl has
//   dE/dT_ka = (L a1_k)_a + (L^T a2_k)_a,   a1_k,b = sum_l G_kl T_lb,  a2_k,b = sum_l G_lk T_lb
//   dE/dr    = sum_k sum_a T_ka (L' a1_k)_a                       (L' = dL/dr: the local integrals' r derivative)
// so one pass over the ten rows k, in plain doubles, yields dE/dr and the 3 x 3 adjoint Y = dE/dR; dE/dx follows
// from the tangents of local_frame (Dual<3>).  L is the local-frame integral matrix, i.e. rotate_xx with R = identity;
// apply_L / apply_LT below write its sparsity out (local packed index: 0 ss | 1,3,6 p_sigma s, p_pi s, p_pi' s |
// 2,4,5,7,8,9 the p p products in SEQM_PP order).
// ---------------------------------------------------------------------------------------------
SEQM_HD void apply_L(const double* ri, const double* v, double* o) {      // o_a = sum_b L_ab v_b
  const double v25 = v[5] + v[9];
  o[0] = ri[0] * v[0] + ri[4] * v[1] + ri[10] * v[2] + ri[11] * v25;
  o[1] = ri[1] * v[0] + ri[5] * v[1] + ri[12] * v[2] + ri[13] * v25;
  o[3] = ri[6] * v[3] + ri[14] * v[4];
  o[6] = ri[6] * v[6] + ri[14] * v[7];
  o[2] = ri[2] * v[0] + ri[7] * v[1] + ri[15] * v[2] + ri[17] * v25;
  o[4] = ri[9] * v[3] + ri[19] * v[4];
  o[5] = ri[3] * v[0] + ri[8] * v[1] + ri[16] * v[2] + ri[18] * v[5] + ri[20] * v[9];
  o[7] = ri[9] * v[6] + ri[19] * v[7];
  o[8] = ri[21] * v[8];
  o[9] = ri[3] * v[0] + ri[8] * v[1] + ri[16] * v[2] + ri[20] * v[5] + ri[18] * v[9];
}
SEQM_HD void apply_LT(const double* ri, const double* v, double* o) {     // o_a = sum_b L_ba v_b
  const double v25 = v[5] + v[9];
  o[0] = ri[0] * v[0] + ri[1] * v[1] + ri[2] * v[2] + ri[3] * v25;
  o[1] = ri[4] * v[0] + ri[5] * v[1] + ri[7] * v[2] + ri[8] * v25;
  o[3] = ri[6] * v[3] + ri[9] * v[4];
  o[6] = ri[6] * v[6] + ri[9] * v[7];
  o[2] = ri[10] * v[0] + ri[12] * v[1] + ri[15] * v[2] + ri[16] * v25;
  o[4] = ri[14] * v[3] + ri[19] * v[4];
  o[5] = ri[11] * v[0] + ri[13] * v[1] + ri[17] * v[2] + ri[18] * v[5] + ri[20] * v[9];
  o[7] = ri[14] * v[6] + ri[19] * v[7];
  o[8] = ri[21] * v[8];
  o[9] = ri[11] * v[0] + ri[13] * v[1] + ri[17] * v[2] + ri[20] * v[5] + ri[18] * v[9];
}

// local-frame images of row k and column k of G:  a1_b = sum_l G(k,l) T_lb,  a2_b = sum_l G(l,k) T_lb
template <class GFun>
SEQM_HD void g_images(const GFun& G, int k, const double R[3][3], const double Tq[6][6], double* a1, double* a2) {
  a1[0] = G(k, 0);
  a2[0] = G(0, k);
  double g1[3], g2[3];
#pragma unroll
  for (int m = 0; m < 3; ++m) { g1[m] = G(k, SEQM_PS(m)); g2[m] = G(SEQM_PS(m), k); }
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    a1[SEQM_PS(a)] = g1[0] * R[0][a] + g1[1] * R[1][a] + g1[2] * R[2][a];
    a2[SEQM_PS(a)] = g2[0] * R[0][a] + g2[1] * R[1][a] + g2[2] * R[2][a];
  }
  double h1[6], h2[6];
#pragma unroll
  for (int q = 0; q < 6; ++q) { h1[q] = G(k, SEQM_PP(q)); h2[q] = G(SEQM_PP(q), k); }
#pragma unroll
  for (int c = 0; c < 6; ++c) {
    double s1 = 0.0, s2 = 0.0;
#pragma unroll
    for (int q = 0; q < 6; ++q) { s1 += h1[q] * Tq[q][c]; s2 += h2[q] * Tq[q][c]; }
    a1[SEQM_PP(c)] = s1;
    a2[SEQM_PP(c)] = s2;
  }
}

// G(k, l) -> upstream dE/dw[k][l];  ri, dri: local integrals and their r derivative;  xij: unit vector
// out: gx[3] = dE/dxij, gr = dE/drij (integral part only)
template <class GFun>
SEQM_HD void rotate_xx_adjoint(const double* ri, const double* dri, const double* xij, const GFun& G, double* gx,
                               double& gr) {
  double R[3][3], Tq[6][6], Y[3][3];
  local_frame(xij, R);
#pragma unroll
  for (int m = 0; m < 3; ++m)
#pragma unroll
    for (int a = 0; a < 3; ++a) Y[m][a] = 0.0;
  build_tq(R, Tq);
  double acc_r = 0.0;
  double a1[10], a2[10], z1[10], z2[10], zr[10];
  {      // k = 0 (ss): T_00 = 1 carries no geometry
    g_images(G, 0, R, Tq, a1, a2);
    apply_L(dri, a1, zr);
    acc_r += zr[0];
  }
#pragma unroll
  for (int m = 0; m < 3; ++m) {      // rows (p_m s): T = R[m][.]
    g_images(G, SEQM_PS(m), R, Tq, a1, a2);
    apply_L(ri, a1, z1);
    apply_LT(ri, a2, z2);
    apply_L(dri, a1, zr);
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      Y[m][a] += z1[SEQM_PS(a)] + z2[SEQM_PS(a)];
      acc_r += R[m][a] * zr[SEQM_PS(a)];
    }
  }
#pragma unroll
  for (int q = 0; q < 6; ++q) {      // rows (p_m p_n): T = Tq[q][.], Tq[q][(a,b)] = R[m][a]R[n][b] + (a != b) R[m][b]R[n][a]
    g_images(G, SEQM_PP(q), R, Tq, a1, a2);
    apply_L(ri, a1, z1);
    apply_LT(ri, a2, z2);
    apply_L(dri, a1, zr);
    const int m = SEQM_QM(q), n = SEQM_QN(q);
#pragma unroll
    for (int c = 0; c < 6; ++c) {
      const int a = SEQM_QM(c), b = SEQM_QN(c);
      const double zc = z1[SEQM_PP(c)] + z2[SEQM_PP(c)];
      acc_r += Tq[q][c] * zr[SEQM_PP(c)];
      Y[m][a] += zc * R[n][b];
      Y[n][b] += zc * R[m][a];
      if (a != b) { Y[m][b] += zc * R[n][a]; Y[n][a] += zc * R[m][b]; }
    }
  }
  gr = acc_r;
  Dual<3> x3[3], R3[3][3];      // tangents of the frame (same branches as the plain evaluation above)
#pragma unroll
  for (int c = 0; c < 3; ++c) { x3[c] = Dual<3>(xij[c]); x3[c].d[c] = 1.0; }
  local_frame(x3, R3);
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    double s = 0.0;
#pragma unroll
    for (int m = 0; m < 3; ++m)
#pragma unroll
      for (int a = 0; a < 3; ++a) s += Y[m][a] * R3[m][a].d[c];
    gx[c] = s;
  }
}

// heavy-hydrogen pair: the ten (mu nu | s s) integrals, sink(k, 0, value)
template <class S, class Sink>
SEQM_HD void rotate_xh(const S* ri, const S R[3][3], Sink& sink) {
  sink(0, 0, ri[0]);
#pragma unroll
  for (int m = 0; m < 3; ++m) sink(SEQM_PS(m), 0, R[m][0] * ri[1]);
#pragma unroll
  for (int q = 0; q < 6; ++q) {
    const int m = SEQM_QM(q), n = SEQM_QN(q);
    // Tq[q][0] = R[m][0]R[n][0];  Tq[q][2] + Tq[q][5] = R[m][1]R[n][1] + R[m][2]R[n][2]
    sink(SEQM_PP(q), 0, ri[2] * (R[m][0] * R[n][0]) + ri[3] * (R[m][1] * R[n][1] + R[m][2] * R[n][2]));
  }
}

// ---------------------------------------------------------------------------------------------
// Slater overlaps.  Auxiliary integrals A_1..A_5(x), B_1..B_5(x) (diat_overlap.py:287-391):
// closed form for |x| > 0.5, four-term series for 1e-6 < |x| <= 0.5, limits below.
// ---------------------------------------------------------------------------------------------
template <class S>
SEQM_HD void aux_A(const S& x, S* A) {
  const S ix = recip_(x);      // one division; the recursions multiply (a few ulp from the divided form)
  A[0] = exp_(-x) * ix;
#pragma unroll
  for (int k = 1; k < 5; ++k) A[k] = A[0] + (double)k * A[k - 1] * ix;
}

template <class S>
SEQM_HD void aux_B(const S& x, S* B) {
  const double ax = fabs(val(x));
  if (ax > 0.5) {
    const S ix = recip_(x);
    const S ep = exp_(x) * ix, em = -(exp_(-x) * ix);
    B[0] = ep + em;
#pragma unroll
    for (int k = 1; k < 5; ++k) B[k] = ((k % 2 == 0) ? ep : -ep) + em + (double)k * B[k - 1] * ix;
  } else if (ax > 1.0e-6) {
    const S x2 = x * x, x3 = x2 * x, x4 = x2 * x2, x5 = x4 * x, x6 = x3 * x3;
    // (coefficients as reciprocal constants: a division by a constant that is not a power of two stays a division)
    B[0] = 2.0 + x2 * (1.0 / 3.0) + x4 * (1.0 / 60.0) + x6 * (1.0 / 2520.0);
    B[1] = -(2.0 / 3.0) * x - x3 * (1.0 / 15.0) - x5 * (1.0 / 420.0);
    B[2] = 2.0 / 3.0 + x2 * (1.0 / 5.0) + x4 * (1.0 / 84.0) + x6 * (1.0 / 3240.0);
    B[3] = -(2.0 / 5.0) * x - x3 * (1.0 / 21.0) - x5 * (1.0 / 540.0);
    B[4] = 2.0 / 5.0 + x2 * (1.0 / 7.0) + x4 * (1.0 / 108.0) + x6 * (1.0 / 3960.0);
  } else {
    B[0] = S(2.0); B[1] = S(0.0); B[2] = S(2.0 / 3.0); B[3] = S(0.0); B[4] = S(2.0 / 5.0);
  }
}

// direction cosines of the pair axis and the p-p rotation block (diat_overlap.py:24-45)
template <class S>
SEQM_HD void overlap_angles(const S* xij, S u[3], S rot[3][3]) {
  const S xy2 = xij[0] * xij[0] + xij[1] * xij[1];
  if (sqrt(val(xy2)) >= 1.0e-10) {
    const S xy = sqrt_(xy2);
    const S ixy = recip_(xy);
    const S ca = xij[0] * ixy, sa = xij[1] * ixy, cb = xij[2], sb = xy;
    u[0] = ca * sb; u[1] = sa * sb; u[2] = cb;
    rot[0][0] = sq_(ca * cb) + sa * sa;
    rot[0][1] = ca * cb * sa * cb - sa * ca;
    rot[0][2] = -(ca * cb * sb);
    rot[1][1] = sq_(sa * cb) + ca * ca;
    rot[1][2] = -(sa * cb * sb);
    rot[2][2] = sb * sb;
  } else {
    const double sg = val(xij[2]) < 0.0 ? -1.0 : (val(xij[2]) > 0.0 ? 1.0 : 0.0);
    u[0] = S(0.0); u[1] = S(0.0); u[2] = S(sg);
    rot[0][0] = S(sg * sg * sg * sg); rot[0][1] = S(0.0); rot[0][2] = S(0.0);
    rot[1][1] = S(sg * sg); rot[1][2] = S(0.0); rot[2][2] = S(0.0);
  }
  rot[1][0] = rot[0][1]; rot[2][0] = rot[0][2]; rot[2][1] = rot[1][2];
}

// the five radial overlap integrals of a heavy-heavy pair: s[0] <s|s>, s[1] <p_sigma|s>, s[2] <s|p_sigma>,
// s[3] <p_sigma|p_sigma>, s[4] <p_pi|p_pi> (diat_overlap.py:96-230)
template <class S>
SEQM_HD void overlap_radial_xx(const S& r, const AtomPar& a, const AtomPar& b, S s[5]) {
  const double SQ3 = 1.7320508075688772;
  S A[5], B[5];
  const S r2 = r * r;
  const S r4 = r2 * r2;
  const S r5 = r4 * r;
  aux_A(0.5 * (a.zs + b.zs) * r, A);
  aux_B(0.5 * (a.zs - b.zs) * r, B);
  s[0] = (b.zs25 * a.zs25 * (1.0 / 48.0)) * r5 * (A[4] * B[0] + B[4] * A[0] - 2.0 * (A[2] * B[2]));
  aux_A(0.5 * (a.zp + b.zs) * r, A);
  aux_B(0.5 * (a.zp - b.zs) * r, B);
  s[1] = (b.zs25 * a.zp25 * (1.0 / (16.0 * SQ3))) * r5 *
         (A[3] * (B[0] - B[2]) - A[1] * (B[2] - B[4]) + B[3] * (A[0] - A[2]) - B[1] * (A[2] - A[4]));
  aux_A(0.5 * (a.zs + b.zp) * r, A);
  aux_B(0.5 * (a.zs - b.zp) * r, B);
  s[2] = (b.zp25 * a.zs25 * (1.0 / (16.0 * SQ3))) * r5 *
         (A[3] * (B[0] - B[2]) - A[1] * (B[2] - B[4]) - B[3] * (A[0] - A[2]) + B[1] * (A[2] - A[4]));
  aux_A(0.5 * (a.zp + b.zp) * r, A);
  aux_B(0.5 * (a.zp - b.zp) * r, B);
  const S pre = (b.zp25 * a.zp25 / 16.0) * r5;
  s[3] = -(pre * (B[2] * (A[4] + A[0]) - A[2] * (B[4] + B[0])));
  s[4] = 0.5 * pre * (A[4] * (B[0] - B[2]) - B[4] * (A[0] - A[2]) - A[2] * B[0] + B[2] * A[0]);
}

// di[mu*4+nu] = <mu on i | nu on j>.  kind: 0 HH, 1 XH (i heavy, j hydrogen), 2 XX.
template <class S>
SEQM_HD void overlap_block(int kind, const S* xij, const S& r, const AtomPar& a, const AtomPar& b, S* di) {
#pragma unroll
  for (int t = 0; t < 16; ++t) di[t] = S(0.0);
  const double SQ3 = 1.7320508075688772;
  S A[5], B[5];
  const S r2 = r * r;
  aux_A(0.5 * (a.zs + b.zs) * r, A);
  aux_B(0.5 * (a.zs - b.zs) * r, B);
  if (kind == 0) {
    di[0] = (a.zs15 * b.zs15 * 0.25) * (r2 * r) * (A[2] * B[0] - B[2] * A[0]);
    return;
  }
  S u[3], rot[3][3];
  overlap_angles(xij, u, rot);
  const S r4 = r2 * r2;
  if (kind == 1) {
    di[0] = (b.zs15 * a.zs25 * (1.0 / (SQ3 * 8.0))) * r4 * (A[3] * B[0] - B[3] * A[0] + A[2] * B[1] - B[2] * A[1]);
    aux_A(0.5 * (a.zp + b.zs) * r, A);
    aux_B(0.5 * (a.zp - b.zs) * r, B);
    const S s211 = (b.zs15 * a.zp25 / 8.0) * r4 * (A[2] * B[0] - B[2] * A[0] + A[3] * B[1] - B[3] * A[1]);
#pragma unroll
    for (int m = 0; m < 3; ++m) di[(1 + m) * 4] = s211 * u[m];
    return;
  }
  S sr[5];
  overlap_radial_xx(r, a, b, sr);
  di[0] = sr[0];
  const S &s211 = sr[1], &s121 = sr[2], &s221 = sr[3], &s222 = sr[4];
#pragma unroll
  for (int m = 0; m < 3; ++m) {
    di[(1 + m) * 4] = s211 * u[m];
    di[1 + m] = -(s121 * u[m]);
#pragma unroll
    for (int n = 0; n < 3; ++n) di[(1 + m) * 4 + 1 + n] = s222 * rot[m][n] - s221 * (u[m] * u[n]);
  }
}

// Reverse mode of the heavy-heavy overlap block: for E = sum F[mu*4+nu] di[mu*4+nu], out[0..2] = dE/dxij, out[3] = dE/dr.
// The block separates into radial integrals times angular factors,
//   E = F00 s0 + s1 c1 - s2 c2 + s4 c3 - s3 c4,   c1 = sum F[p_m][s] u_m,  c2 = sum F[s][p_m] u_m,
//                                                  c3 = sum F[p_m][p_n] rot_mn,  c4 = sum F[p_m][p_n] u_m u_n,
// so the radial part (four pairs of auxiliary integrals: the expensive half) carries ONE tangent (d/dr) and the cheap
// angular part three, instead of four tangents through both and 16 dual-number results.
SEQM_HD void overlap_xx_adjoint(const double* xij, double r, const AtomPar& a, const AtomPar& b, const double* F,
                                double* out) {
  Dual<3> x3[3], u[3], rot[3][3];
#pragma unroll
  for (int c = 0; c < 3; ++c) { x3[c] = Dual<3>(xij[c]); x3[c].d[c] = 1.0; }
  overlap_angles(x3, u, rot);
  Dual<3> c1(0.0), c2(0.0), c3(0.0), c4(0.0);
#pragma unroll
  for (int m = 0; m < 3; ++m) {
    c1 = c1 + F[(1 + m) * 4] * u[m];
    c2 = c2 + F[1 + m] * u[m];
#pragma unroll
    for (int n = 0; n < 3; ++n) {
      c3 = c3 + F[(1 + m) * 4 + 1 + n] * rot[m][n];
      c4 = c4 + F[(1 + m) * 4 + 1 + n] * (u[m] * u[n]);
    }
  }
  Dual<1> r1(r), s[5];
  r1.d[0] = 1.0;
  overlap_radial_xx(r1, a, b, s);
#pragma unroll
  for (int c = 0; c < 3; ++c) out[c] = s[1].v * c1.d[c] - s[2].v * c2.d[c] + s[4].v * c3.d[c] - s[3].v * c4.d[c];
  out[3] = F[0] * s[0].d[0] + s[1].d[0] * c1.v - s[2].d[0] * c2.v + s[4].d[0] * c3.v - s[3].d[0] * c4.v;
}

// core-core repulsion of one pair (energy.py:38-78); gam = (ss|ss); ngauss = 0 (MNDO), 4 (AM1), 2 (PM3)
template <class S>
SEQM_HD S core_repulsion(int ni, int nj, const S& r, const S& gam, const AtomPar& a, const AtomPar& b, int ngauss) {
  const S Rang = r * SEQM_A0;
  S ea = exp_(-(a.alpha * Rang));
  if ((ni == 7 || ni == 8) && nj == 1) ea = ea * Rang;
  const S eb = exp_(-(b.alpha * Rang));
  const double zz = a.tore * b.tore;
  S E = zz * gam * (1.0 + ea + eb);
  if (ngauss > 0) {
    S g(0.0);
    for (int k = 0; k < ngauss; ++k) {
      g = g + a.gK[k] * exp_(-(a.gL[k] * sq_(Rang - a.gM[k])));
      g = g + b.gK[k] * exp_(-(b.gL[k] * sq_(Rang - b.gM[k])));
    }
    E = E + (zz / Rang) * g;
  }
  return E;
}

// pair kind from the atomic numbers: 0 HH, 1 XH, 2 XX (atoms are sorted so that Z_i >= Z_j)
SEQM_HD int pair_kind(int ni, int nj) { return (ni > 1 ? 1 : 0) + (nj > 1 ? 1 : 0); }

}  // namespace seqm
