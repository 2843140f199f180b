// TEST INFRASTRUCTURE ONLY — never part of the product library.
// Compiles the templated per-pair mathematics of pyseqm_b200/csrc/pair_math.cuh for the HOST so the
// formulas (and their dual-number derivatives) can be checked against the oracle on a machine without a
// GPU.  The product reaches the same functions only from CUDA kernels.
#include "../../pyseqm_b200/csrc/pair_math.cuh"

using namespace seqm;

struct ArrSink {
  double* w;
  void operator()(int k, int l, double v) const { w[k * 10 + l] = v; }
};
struct DualArrSink {
  double* w;   // [100][5]
  void operator()(int k, int l, const Dual<4>& v) const {
    double* o = w + (k * 10 + l) * 5;
    o[0] = v.v;
    for (int q = 0; q < 4; ++q) o[1 + q] = v.d[q];
  }
};

extern "C" {

int hc_atompar_doubles() { return (int)(sizeof(AtomPar) / sizeof(double)); }

// forward values: w[100], di[16], enuc
void hc_pair(int ni, int nj, const double* x, double r, const double* apar, const double* bpar, int ngauss,
             double* w, double* di, double* enuc) {
  const AtomPar& a = *(const AtomPar*)apar;
  const AtomPar& b = *(const AtomPar*)bpar;
  const int kind = pair_kind(ni, nj);
  for (int q = 0; q < 100; ++q) w[q] = 0.0;
  double R[3][3];
  ArrSink sink{w};
  if (kind == 2) {
    double ri[22];
    repp_xx(r, a.mp, b.mp, ri);
    local_frame(x, R);
    rotate_xx(ri, R, sink);
  } else if (kind == 1) {
    double ri[4];
    repp_xh(r, a.mp, b.mp, ri);
    local_frame(x, R);
    rotate_xh(ri, R, sink);
  } else {
    double ri[1];
    repp_hh(r, a.mp, b.mp, ri);
    w[0] = ri[0];
  }
  overlap_block(kind, x, r, a, b, di);
  *enuc = core_repulsion(ni, nj, r, w[0], a, b, ngauss);
}

// duals: w[100][5], di[16][5], enuc[5]  (value, d/dx1, d/dx2, d/dx3, d/dr)
void hc_pair_dual(int ni, int nj, const double* xv, double rv, const double* apar, const double* bpar, int ngauss,
                  double* w, double* di, double* enuc) {
  typedef Dual<4> D4;
  const AtomPar& a = *(const AtomPar*)apar;
  const AtomPar& b = *(const AtomPar*)bpar;
  const int kind = pair_kind(ni, nj);
  for (int q = 0; q < 500; ++q) w[q] = 0.0;
  D4 x[3], r(rv);
  for (int c = 0; c < 3; ++c) { x[c] = D4(xv[c]); x[c].d[c] = 1.0; }
  r.d[3] = 1.0;
  D4 R[3][3];
  DualArrSink sink{w};
  D4 gam;
  if (kind == 2) {
    D4 ri[22];
    repp_xx(r, a.mp, b.mp, ri);
    local_frame(x, R);
    rotate_xx(ri, R, sink);
    gam = ri[0];
  } else if (kind == 1) {
    D4 ri[4];
    repp_xh(r, a.mp, b.mp, ri);
    local_frame(x, R);
    rotate_xh(ri, R, sink);
    gam = ri[0];
  } else {
    D4 ri[1];
    repp_hh(r, a.mp, b.mp, ri);
    sink(0, 0, ri[0]);
    gam = ri[0];
  }
  D4 d[16];
  overlap_block(kind, x, r, a, b, d);
  for (int q = 0; q < 16; ++q) { di[q * 5] = d[q].v; for (int c = 0; c < 4; ++c) di[q * 5 + 1 + c] = d[q].d[c]; }
  D4 e = core_repulsion(ni, nj, r, gam, a, b, ngauss);
  enuc[0] = e.v;
  for (int c = 0; c < 4; ++c) enuc[1 + c] = e.d[c];
}

// reverse-mode adjoint of the heavy-heavy rotation (rotate_xx_adjoint): out[4] = dE/dx1, dx2, dx3, dr for E = sum G.w
void hc_rotate_xx_adjoint(const double* x, double rv, const double* apar, const double* bpar, const double* G, double* out) {
  const AtomPar& a = *(const AtomPar*)apar;
  const AtomPar& b = *(const AtomPar*)bpar;
  Dual<1> r(rv);
  r.d[0] = 1.0;
  Dual<1> ri1[22];
  repp_xx(r, a.mp, b.mp, ri1);
  double ri[22], dri[22];
  for (int q = 0; q < 22; ++q) { ri[q] = ri1[q].v; dri[q] = ri1[q].d[0]; }
  auto g = [G](int k, int l) { return G[k * 10 + l]; };
  rotate_xx_adjoint(ri, dri, x, g, out, out[3]);
}

// reverse-mode adjoint of the heavy-heavy overlap block (overlap_xx_adjoint): out[4] for E = sum F.di
void hc_overlap_xx_adjoint(const double* x, double rv, const double* apar, const double* bpar, const double* F, double* out) {
  overlap_xx_adjoint(x, rv, *(const AtomPar*)apar, *(const AtomPar*)bpar, F, out);
}
}
