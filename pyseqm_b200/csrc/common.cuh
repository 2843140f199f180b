// Shared definitions for the seqm_b200 kernels: constants of the reference (MOPAC7 units), a small
// forward-mode dual number used by the force adjoint, and launch/error helpers.
#pragma once
#include <cmath>
#include <cstdint>

#if defined(__CUDACC__)
#define SEQM_HD __host__ __device__ __forceinline__
#else
#define SEQM_HD inline
#endif

// reference: seqm/seqm_functions/constants.py:3-16 (non-CODATA values, reproduced exactly)
#define SEQM_EV 27.21
#define SEQM_A0 0.529167
#define SEQM_OVERLAP_CUTOFF 40.0

namespace seqm {

// ---------------------------------------------------------------------------------------------
// Dual<N>: value + N tangents.  The pair functions are templated on the scalar so the forward
// kernel (double) and the adjoint kernel (Dual<4>: d/dx1, d/dx2, d/dx3, d/dr) share one source and
// take the same branches.
// ---------------------------------------------------------------------------------------------
template <int N>
struct Dual {
  double v;
  double d[N];
  SEQM_HD Dual() {}
  SEQM_HD Dual(double a) : v(a) {
#pragma unroll
    for (int i = 0; i < N; ++i) d[i] = 0.0;
  }
};

template <int N> SEQM_HD Dual<N> operator+(const Dual<N>& a, const Dual<N>& b) {
  Dual<N> r; r.v = a.v + b.v;
#pragma unroll
  for (int i = 0; i < N; ++i) r.d[i] = a.d[i] + b.d[i];
  return r;
}
template <int N> SEQM_HD Dual<N> operator-(const Dual<N>& a, const Dual<N>& b) {
  Dual<N> r; r.v = a.v - b.v;
#pragma unroll
  for (int i = 0; i < N; ++i) r.d[i] = a.d[i] - b.d[i];
  return r;
}
template <int N> SEQM_HD Dual<N> operator-(const Dual<N>& a) {
  Dual<N> r; r.v = -a.v;
#pragma unroll
  for (int i = 0; i < N; ++i) r.d[i] = -a.d[i];
  return r;
}
template <int N> SEQM_HD Dual<N> operator*(const Dual<N>& a, const Dual<N>& b) {
  Dual<N> r; r.v = a.v * b.v;
#pragma unroll
  for (int i = 0; i < N; ++i) r.d[i] = a.v * b.d[i] + a.d[i] * b.v;
  return r;
}
template <int N> SEQM_HD Dual<N> operator/(const Dual<N>& a, const Dual<N>& b) {
  Dual<N> r; const double ib = 1.0 / b.v; r.v = a.v * ib;
#pragma unroll
  for (int i = 0; i < N; ++i) r.d[i] = (a.d[i] - r.v * b.d[i]) * ib;
  return r;
}
template <int N> SEQM_HD Dual<N> operator+(const Dual<N>& a, double b) { Dual<N> r = a; r.v += b; return r; }
template <int N> SEQM_HD Dual<N> operator+(double b, const Dual<N>& a) { Dual<N> r = a; r.v += b; return r; }
template <int N> SEQM_HD Dual<N> operator-(const Dual<N>& a, double b) { Dual<N> r = a; r.v -= b; return r; }
template <int N> SEQM_HD Dual<N> operator-(double b, const Dual<N>& a) { Dual<N> r = -a; r.v += b; return r; }
template <int N> SEQM_HD Dual<N> operator*(const Dual<N>& a, double b) {
  Dual<N> r; r.v = a.v * b;
#pragma unroll
  for (int i = 0; i < N; ++i) r.d[i] = a.d[i] * b;
  return r;
}
template <int N> SEQM_HD Dual<N> operator*(double b, const Dual<N>& a) { return a * b; }
template <int N> SEQM_HD Dual<N> operator/(const Dual<N>& a, double b) { return a * (1.0 / b); }
template <int N> SEQM_HD Dual<N> operator/(double a, const Dual<N>& b) {
  Dual<N> r; const double ib = 1.0 / b.v; r.v = a * ib; const double f = -r.v * ib;
#pragma unroll
  for (int i = 0; i < N; ++i) r.d[i] = f * b.d[i];
  return r;
}
template <int N> SEQM_HD Dual<N>& operator+=(Dual<N>& a, const Dual<N>& b) { a = a + b; return a; }
template <int N> SEQM_HD Dual<N>& operator-=(Dual<N>& a, const Dual<N>& b) { a = a - b; return a; }

SEQM_HD double val(double a) { return a; }
template <int N> SEQM_HD double val(const Dual<N>& a) { return a.v; }

SEQM_HD double sq_(double a) { return a * a; }
template <int N> SEQM_HD Dual<N> sq_(const Dual<N>& a) { return a * a; }

SEQM_HD double recip_(double a) { return 1.0 / a; }
template <int N> SEQM_HD Dual<N> recip_(const Dual<N>& a) {
  Dual<N> r; r.v = 1.0 / a.v; const double f = -(r.v * r.v);
#pragma unroll
  for (int i = 0; i < N; ++i) r.d[i] = f * a.d[i];
  return r;
}
SEQM_HD double sqrt_(double a) { return sqrt(a); }
template <int N> SEQM_HD Dual<N> sqrt_(const Dual<N>& a) {
  Dual<N> r; r.v = sqrt(a.v); const double f = 0.5 / r.v;
#pragma unroll
  for (int i = 0; i < N; ++i) r.d[i] = f * a.d[i];
  return r;
}
SEQM_HD double exp_(double a) { return exp(a); }
template <int N> SEQM_HD Dual<N> exp_(const Dual<N>& a) {
  Dual<N> r; r.v = exp(a.v);
#pragma unroll
  for (int i = 0; i < N; ++i) r.d[i] = r.v * a.d[i];
  return r;
}
SEQM_HD double abs_(double a) { return fabs(a); }
template <int N> SEQM_HD Dual<N> abs_(const Dual<N>& a) { return a.v < 0.0 ? -a : a; }

// 1/sqrt(u^2 + c), c a plain number (depends on atomic parameters only)
// (device: rsqrt, 1 ulp, instead of an IEEE square root followed by an IEEE division with their slow-path calls - the
// heavy-heavy integrals take 69 of these per pair, a third of the instructions of the pair pass)
// Branch free: the hardware seed (MUFU.RSQ64H, ~2^-22) and ONE third-order step y (1 + e/2 + 3 e^2/8), e = 1 - a y^2,
// which leaves e^3 ~ 2^-60 below the rounding of the five operations.  The arguments here are u^2 + c with c a squared
// sum of charge-separation lengths: normal numbers, so the special-case path of rsqrt() (a convergence barrier and a
// call per use, which also keeps the 69 independent chains from being interleaved) is not needed.
SEQM_HD double rsqrt_(double a) {
#ifdef __CUDA_ARCH__
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(a));
  const double e = fma(-(a * y), y, 1.0);
  return fma(y * e, fma(0.375, e, 0.5), y);
#else
  return 1.0 / sqrt(a);
#endif
}
SEQM_HD double inv_dist(double u, double c) { return rsqrt_(u * u + c); }
template <int N> SEQM_HD Dual<N> inv_dist(const Dual<N>& u, double c) {
  Dual<N> r; r.v = rsqrt_(u.v * u.v + c); const double f = -u.v * r.v * r.v * r.v;
#pragma unroll
  for (int i = 0; i < N; ++i) r.d[i] = f * u.d[i];
  return r;
}

// per-atom point-charge multipole parameters (reference: two_elec_two_center_int.py:22-43)
struct AtomMultipole {
  double dd, qq, rho0, rho1, rho2;
};

}  // namespace seqm
