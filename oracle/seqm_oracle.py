"""
ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product.

A CPU (torch fp64 tensors used as a plain array library, plus autograd for the
coordinate derivative exactly as the reference's `Force` does) restatement of
PYSEQM's batched SCF hot path.  Only `tests/`, `__graft_entry__.smoke()` and
`bench.py`'s cpu_baseline / `--impl reference` legs may import this file; the
product (`pyseqm_b200/`) never does and fails loudly without its CUDA library.

Parity pin: this restatement is checked against the UNMODIFIED reference
imported from /root/reference in the build container by
`tools/make_golden.py`, which also writes the committed fixtures under
`tests/golden/`; `tests/test_oracle_golden.py` re-checks the oracle against
those fixtures without needing the reference (the GPU box has no reference).

Each function cites the reference file:line it restates (paths relative to the
reference root).  The restatement is written independently: e.g. the local
two-centre integrals are evaluated from the Dewar-Thiel point-charge multipole
model generically (seqm/seqm_functions/two_elec_two_center_int_local_frame.py
lists the same sums term by term) and the molecular-frame rotation is a plain
tensor transformation T·L·Tᵀ (two_elec_two_center_int.py:304-702 writes the
100 products out).
"""
import math
import os

import torch

# --- constants (seqm/seqm_functions/constants.py:3-16,55-113) ---------------
EV = 27.21
A0 = 0.529167
EV_KCALPMOL = 23.061
OVERLAP_CUTOFF = 40.0
MAX_ITER = 1000

_T = lambda v: torch.tensor(v, dtype=torch.float64)
TORE = _T([0, 1, 0, 1, 2, 3, 4, 5, 6, 7, 0, 1, 2, 3, 4, 5, 6, 7, 0])
QN = _T([0, 1, 0, 2, 2, 2, 2, 2, 2, 2, 0, 3, 3, 3, 3, 3, 3, 3, 0])
USSC = _T([0, 1, 0, 1, 2, 2, 2, 2, 2, 2, 0, 1, 2, 2, 2, 2, 2, 2, 0])
UPPC = _T([0, 0, 0, 0, 0, 1, 2, 3, 4, 5, 6, 0, 0, 1, 2, 3, 4, 5, 6])
GSSC = _T([0, 0, 0, 0, 1, 1, 1, 1, 1, 1, 0, 0, 1, 1, 1, 1, 1, 1, 0])
GSPC = _T([0, 0, 0, 0, 0, 2, 4, 6, 8, 10, 0, 0, 0, 2, 4, 6, 8, 10, 0])
HSPC = _T([0, 0, 0, 0, 0, -1, -2, -3, -4, -5, 0, 0, 0, -1, -2, -3, -4, -5, 0])
GP2C = _T([0, 0, 0, 0, 0, 0, 1.5, 4.5, 6.5, 10, 0, 0, 0, 0, 1.5, 4.5, 6.5, 10, 0])
GPPC = _T([0, 0, 0, 0, 0, 0, -0.5, -1.5, -0.5, 0, 0, 0, 0, 0, -0.5, -1.5, -0.5, 0, 0])
EHEAT = _T([0.0, 52.102, 0.0, 38.410, 76.960, 135.700, 170.890, 113.000, 59.559, 18.890, 0.0,
            25.850, 35.000, 79.490, 108.390, 75.570, 66.400, 28.990, 0.0]) / EV_KCALPMOL
MASS = _T([0.0, 1.00790, 4.00260, 6.94000, 9.01218, 10.81000, 12.01100, 14.00670, 15.99940,
           18.99840, 20.17900, 22.98977, 24.30500, 26.98154, 28.08550, 30.97376, 32.06000,
           35.45300, 39.94800])

BASE_PARAMS = ['U_ss', 'U_pp', 'zeta_s', 'zeta_p', 'beta_s', 'beta_p',
               'g_ss', 'g_sp', 'g_pp', 'g_p2', 'h_sp', 'alpha']
N_GAUSS = {'MNDO': 0, 'AM1': 4, 'PM3': 2}

_PARAM_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', 'pyseqm_b200', 'params')


def load_params(method, param_dir=None):
    """Element parameter table {name: tensor[19]} (seqm/seqm_functions/parameters.py:4-31).

    Gaussians come back stacked: 'K','L','M' -> tensor[19, ng] (basics.py:293-320)."""
    fn = os.path.join(param_dir or _PARAM_DIR, 'parameters_%s_MOPAC.csv' % method)
    with open(fn) as f:
        header = [h.strip() for h in f.readline().split(',')]
        table = {}
        for line in f:
            t = [x.strip() for x in line.split(',')]
            if t and t[0]:
                table[int(t[0])] = t
    ng = N_GAUSS[method]
    names = list(BASE_PARAMS)
    for g in range(1, ng + 1):
        names += ['Gaussian%d_K' % g, 'Gaussian%d_L' % g, 'Gaussian%d_M' % g]
    out = {}
    for nm in names:
        col = header.index(nm)
        v = torch.zeros(19, dtype=torch.float64)
        for z, t in table.items():
            if z <= 18:
                v[z] = float(t[col])
        out[nm] = v
    for c in 'KLM':
        if ng:
            out[c] = torch.stack([out['Gaussian%d_%s' % (g, c)] for g in range(1, ng + 1)], dim=1)
    return out


# --- Parser (seqm/basics.py:43-118) -----------------------------------------
def parse(species, coordinates, cutoff=1.0e10, charges=None):
    """species[B,S] int64 (descending Z, 0 padded), coordinates[B,S,3] (Å).

    Returns a dict with the 16 reference outputs.  Pairs are ordered (mol, a, b), a<b."""
    B, S = species.shape
    real = species > 0
    flat_real = real.reshape(-1)
    real_atoms = torch.nonzero(flat_real).reshape(-1)
    inv = torch.zeros(B * S, dtype=torch.int64)
    inv[real_atoms] = torch.arange(real_atoms.numel())
    Z = species.reshape(-1)[real_atoms]
    nHeavy = (species > 1).sum(1)
    nHydro = (species == 1).sum(1)
    nel = TORE[species].sum(1).to(torch.int64)
    if charges is not None:
        nel = nel - charges.reshape(-1).to(torch.int64)
    if ((nel % 2) == 1).any():
        raise ValueError("Only closed shell system (with even number of electrons) are supported")
    nocc = torch.div(nel, 2, rounding_mode='floor')
    mol_of = torch.arange(B).repeat_interleave(S)
    slot_of = torch.arange(S).repeat(B)
    maskd = (mol_of * S * S + slot_of * (S + 1))[real_atoms]
    atom_molid = mol_of[real_atoms]
    # all a<b slot pairs of one molecule, row-major == (a, b) lexicographic
    a, b = torch.triu_indices(S, S, offset=1)
    d = coordinates[:, b, :] - coordinates[:, a, :]            # R_b - R_a   [B, npair, 3]
    d2 = (d ** 2).sum(dim=2)
    d2 = torch.where(d2 == 0.0, torch.full_like(d2, 1.0e-4), d2)   # basics.py:93
    dist = torch.sqrt(d2)
    keep = real[:, a] & real[:, b] & (dist < cutoff)
    keep_f = keep.reshape(-1)
    dvec = d.reshape(-1, 3)[keep_f]
    dist = dist.reshape(-1)[keep_f]
    mol = torch.arange(B).unsqueeze(1).expand(B, a.numel()).reshape(-1)[keep_f]
    ga = (mol * S + a.repeat(B)[keep_f])
    gb = (mol * S + b.repeat(B)[keep_f])
    idxi, idxj = inv[ga], inv[gb]
    return dict(nmol=B, molsize=S, nHeavy=nHeavy, nHydro=nHydro, nocc=nocc, Z=Z, maskd=maskd,
                atom_molid=atom_molid, mask=ga * S + gb % S, pair_molid=mol,
                ni=Z[idxi], nj=Z[idxj], idxi=idxi, idxj=idxj,
                xij=dvec / dist.unsqueeze(1), rij=dist * (1.0 / A0))


# --- per-atom multipole parameters (cal_par.py:8-20,30-90,118-173) ----------
def _secant(fun, target, x1):
    x2 = x1 + 0.04
    for _ in range(19):
        f1, f2 = fun(x1), fun(x2)
        x3 = torch.where((f2 - f1).abs() > 1.0e-16, x1 + (x2 - x1) * (target - f1) / (f2 - f1), x2)
        x1, x2 = x2, x3
        if (fun(x2) - target).abs().max() < 1.0e-16:
            break
    return x2


def atom_multipoles(Z, zs, zp, gss, gpp, gp2, hsp):
    """dd, qq, rho0, rho1, rho2 per atom (two_elec_two_center_int.py:22-43)."""
    heavy = Z > 2
    dd = torch.zeros_like(zs)
    qq = torch.zeros_like(zs)
    rho0 = torch.zeros_like(zs)
    rho1 = torch.zeros_like(zs)
    rho2 = torch.zeros_like(zs)
    ok = (Z == 1) | heavy
    rho0[ok] = 0.5 * EV / gss[ok]
    if heavy.any():
        n = QN[Z][heavy]
        s, p = zs[heavy], zp[heavy]
        D1 = (2.0 * n + 1.0) * (4.0 * s * p) ** (n + 0.5) / (s + p) ** (2.0 * n + 2.0) / math.sqrt(3.0)
        D2 = torch.sqrt((4.0 * n ** 2 + 6.0 * n + 2.0) / 20.0) / p
        dd[heavy], qq[heavy] = D1, D2
        h1 = hsp[heavy] / EV
        g1 = (h1.abs() / D1 ** 2) ** (1.0 / 3.0)
        g1 = torch.where(h1 < 0, -g1, g1)
        d = _secant(lambda x: 0.5 * x - 0.5 / torch.sqrt(4.0 * D1 ** 2 + 1.0 / x ** 2), h1, g1)
        rho1[heavy] = 0.5 / d
        h2 = 0.5 * (gpp[heavy] - gp2[heavy]) / EV
        g2 = (h2.abs() / 3.0 / D2 ** 4) ** 0.2
        g2 = torch.where(h2 < 0, -g2, g2)
        q = _secant(lambda x: 0.25 * x - 0.5 / torch.sqrt(4.0 * D2 ** 2 + 1.0 / x ** 2)
                    + 0.25 / torch.sqrt(8.0 * D2 ** 2 + 1.0 / x ** 2), h2, g2)
        rho2[heavy] = 0.5 / q
    return dd, qq, rho0, rho1, rho2


# --- local-frame two-centre integrals ---------------------------------------
# Dewar-Thiel multipoles as point charges: (coefficient, unit offset (x,y,z), separation kind
# [0 none, 1 dd, 2 qq], additive-term kind [0 rho0, 1 rho1, 2 rho2]).  Local orbital order is
# (s, sigma, pi, pi'), sigma along the axis j->i; pair index k = 0 ss, 1 sigma-s, 2 sigma-sigma,
# 3 pi-s, 4 pi-sigma, 5 pi-pi, 6 pi'-s, 7 pi'-sigma, 8 pi'-pi, 9 pi'-pi'.
def _dip(ax):
    u = [0, 0, 0]
    u[ax] = 1
    v = [0, 0, 0]
    v[ax] = -1
    return [(0.5, tuple(u), 1, 1), (-0.5, tuple(v), 1, 1)]


def _lin_quad(ax):
    u = [0, 0, 0]
    u[ax] = 2
    v = [0, 0, 0]
    v[ax] = -2
    return [(1.0, (0, 0, 0), 0, 0), (0.25, tuple(u), 2, 2), (0.25, tuple(v), 2, 2), (-0.5, (0, 0, 0), 0, 2)]


def _sq_quad(ax, bx):
    out = []
    for sa in (1, -1):
        for sb in (1, -1):
            u = [0, 0, 0]
            u[ax] = sa
            u[bx] = sb
            out.append((0.25 * sa * sb, tuple(u), 2, 2))
    return out


_X, _Y, _ZAX = 0, 1, 2     # pi, pi', sigma directions of the local frame
_MULTIPOLE = {0: [(1.0, (0, 0, 0), 0, 0)], 1: _dip(_ZAX), 2: _lin_quad(_ZAX), 3: _dip(_X),
              4: _sq_quad(_X, _ZAX), 5: _lin_quad(_X), 6: _dip(_Y), 7: _sq_quad(_Y, _ZAX),
              9: _lin_quad(_Y)}
# the 22 unique integrals in MOPAC repp order (two_elec_two_center_int_local_frame.py:39-43)
_RI_KL = [(0, 0), (1, 0), (2, 0), (5, 0), (0, 1), (1, 1), (3, 3), (2, 1), (5, 1), (4, 3), (0, 2),
          (0, 5), (1, 2), (1, 5), (3, 4), (2, 2), (5, 2), (2, 5), (5, 5), (4, 4), (5, 9)]
# where each unique integral sits in the local 10x10 tensor (pi <-> pi' equivalence)
_RI_PLACES = [[(0, 0)], [(1, 0)], [(2, 0)], [(5, 0), (9, 0)], [(0, 1)], [(1, 1)], [(3, 3), (6, 6)],
              [(2, 1)], [(5, 1), (9, 1)], [(4, 3), (7, 6)], [(0, 2)], [(0, 5), (0, 9)], [(1, 2)],
              [(1, 5), (1, 9)], [(3, 4), (6, 7)], [(2, 2)], [(5, 2), (9, 2)], [(2, 5), (2, 9)],
              [(5, 5), (9, 9)], [(4, 4), (7, 7)], [(5, 9), (9, 5)], [(8, 8)]]


def _charges(k, dd, qq, rhos):
    out = []
    for c, u, sk, rk in _MULTIPOLE[k]:
        sep = (None, dd, qq)[sk]
        pos = [0.0 if (ui == 0 or sep is None) else ui * sep for ui in u]
        out.append((c, pos, rhos[rk]))
    return out


def local_integrals(r, da, db, qa, qb, rho_a, rho_b):
    """ri[P,22] in eV (two_elec_two_center_int_local_frame.py:113-261) for every pair treated as
    heavy-heavy; the caller masks orbitals that do not exist on hydrogen."""
    ri = []
    for k, l in _RI_KL:
        acc = 0.0
        for ca, pa, ra in _charges(k, da, qa, rho_a):
            for cb, pb, rb in _charges(l, db, qb, rho_b):
                dx, dy = pa[0] - pb[0], pa[1] - pb[1]
                dz = r + pa[2] - pb[2]
                acc = acc + (ca * cb * EV) / torch.sqrt(dx * dx + dy * dy + dz * dz + (ra + rb) ** 2)
        ri.append(acc)
    ri.append(0.5 * (ri[18] - ri[20]))      # (pi'pi|pi'pi) by rotational invariance (:261)
    return torch.stack(ri, dim=1)


_IND = [[0, 1, 3, 6], [1, 2, 4, 7], [3, 4, 5, 8], [6, 7, 8, 9]]     # fock.py:120-123
_PAIRS10 = [(0, 0), (1, 0), (1, 1), (2, 0), (2, 1), (2, 2), (3, 0), (3, 1), (3, 2), (3, 3)]


def local_frame(xij):
    """Axes X (sigma), Y, Z of MOPAC rotate (two_elec_two_center_int.py:210-265)."""
    X = -xij
    x3 = X[:, 2]
    inside = x3.abs() < 1.0
    z3 = torch.zeros_like(x3)
    z3[inside] = torch.sqrt(1.0 - x3[inside] ** 2)
    gen = z3 > 1.0e-5
    z1 = torch.ones_like(x3)
    z2 = torch.zeros_like(x3)
    y1 = torch.zeros_like(x3)
    y2 = torch.ones_like(x3)
    a = 1.0 / z3[gen]
    z1[gen] = -a * X[gen, 0] * x3[gen]
    z2[gen] = -a * X[gen, 1] * x3[gen]
    sgn = torch.where(X[gen, 0] >= 0.0, -torch.ones_like(a), torch.ones_like(a))
    y1[gen] = sgn * a * X[gen, 1]
    y2[gen] = (a * X[gen, 0]).abs()
    Y = torch.stack((y1, y2, torch.zeros_like(y1)), dim=1)
    Zv = torch.stack((z1, z2, z3), dim=1)
    return X, Y, Zv


def two_center_integrals(ni, nj, xij, rij, dd, qq, rho0, rho1, rho2, idxi, idxj):
    """w[P,10,10], e1b[P,4,4], e2a[P,4,4] (two_elec_two_center_int.py:56-878)."""
    P = rij.shape[0]
    ri = local_integrals(rij, dd[idxi], dd[idxj], qq[idxi], qq[idxj],
                         (rho0[idxi], rho1[idxi], rho2[idxi]), (rho0[idxj], rho1[idxj], rho2[idxj]))
    L = torch.zeros(P, 10, 10, dtype=rij.dtype)
    for n, places in enumerate(_RI_PLACES):
        for k, l in places:
            L[:, k, l] = ri[:, n]
    X, Y, Zv = local_frame(xij)
    O = torch.zeros(P, 4, 4, dtype=rij.dtype)          # molecular orbital mu -> local orbital a
    O[:, 0, 0] = 1.0
    O[:, 1:, 1] = X
    O[:, 1:, 2] = Y
    O[:, 1:, 3] = Zv
    T = torch.zeros(P, 10, 10, dtype=rij.dtype)
    for k, (m, n) in enumerate(_PAIRS10):
        for c, (a, b) in enumerate(_PAIRS10):
            t = O[:, m, a] * O[:, n, b]
            if a != b:
                t = t + O[:, m, b] * O[:, n, a]
            T[:, k, c] = t
    w = torch.einsum('pka,pab,plb->pkl', T, L, T)
    vi = torch.zeros(P, 10, dtype=rij.dtype)
    vj = torch.zeros(P, 10, dtype=rij.dtype)
    vi[:, 0] = 1.0
    vj[:, 0] = 1.0
    vi[ni > 1] = 1.0
    vj[nj > 1] = 1.0
    w = w * vi.unsqueeze(2) * vj.unsqueeze(1)
    e1b = torch.zeros(P, 4, 4, dtype=rij.dtype)
    e2a = torch.zeros(P, 4, 4, dtype=rij.dtype)
    for k, (m, n) in enumerate(_PAIRS10):       # upper triangle [n, m], n <= m
        e1b[:, n, m] = -TORE[nj] * w[:, k, 0]
        e2a[:, n, m] = -TORE[ni] * w[:, 0, k]
    return w, e1b, e2a


# --- STO overlap (diat_overlap.py:3-393) ------------------------------------
def _aux_A(x):
    """A_1..A_5 (diat_overlap.py:287-296)."""
    a = [torch.exp(-x) / x]
    for k in range(1, 5):
        a.append(a[0] + k * a[-1] / x)
    return a


def _aux_B(x):
    """B_1..B_5: closed form |x|>0.5, 4-term series 1e-6<|x|<=0.5, limits below (:319-391)."""
    big = x.abs() > 0.5
    mid = (x.abs() <= 0.5) & (x.abs() > 1.0e-6)
    xs = torch.where(big, x, torch.ones_like(x))
    ep, em = torch.exp(xs) / xs, -torch.exp(-xs) / xs
    closed = [ep + em]
    for k in range(1, 5):
        closed.append((ep if k % 2 == 0 else -ep) + em + k * closed[-1] / xs)
    x2, x3, x4, x5, x6 = x ** 2, x ** 3, x ** 4, x ** 5, x ** 6
    series = [2.0 + x2 / 3.0 + x4 / 60.0 + x6 / 2520.0,
              -2.0 / 3.0 * x - x3 / 15.0 - x5 / 420.0,
              2.0 / 3.0 + x2 / 5.0 + x4 / 84.0 + x6 / 3240.0,
              -2.0 / 5.0 * x - x3 / 21.0 - x5 / 540.0,
              2.0 / 5.0 + x2 / 7.0 + x4 / 108.0 + x6 / 3960.0]
    limit = [2.0, 0.0, 2.0 / 3.0, 0.0, 2.0 / 5.0]
    return [torch.where(big, c, torch.where(mid, s, torch.full_like(x, l)))
            for c, s, l in zip(closed, series, limit)]


def _AB(r, z1, z2):
    return _aux_A(0.5 * r * (z1 + z2)), _aux_B(0.5 * r * (z1 - z2))


def overlap(ni, nj, xij, rij, zeta_a, zeta_b):
    """di[P,4,4]: di[mu on i, nu on j] (diat_overlap.py:3-246)."""
    P = rij.shape[0]
    qi, qj = QN[ni], QN[nj]
    hh = (qi == 1) & (qj == 1)
    xh = (qi == 2) & (qj == 1)
    xx = (qi == 2) & (qj == 2)
    if not (hh | xh | xx).all():
        raise ValueError("overlap: unsupported element (valence shell must be n=1 or n=2)")
    zas, zap, zbs, zbp = zeta_a[:, 0], zeta_a[:, 1], zeta_b[:, 0], zeta_b[:, 1]
    zero = torch.zeros_like(rij)
    sq3 = math.sqrt(3.0)
    r = rij
    A, B = _AB(r, zas, zbs)
    s_hh = (zas * zbs * r ** 2) ** 1.5 * (A[2] * B[0] - B[2] * A[0]) / 4.0
    s_xh = zbs ** 1.5 * zas ** 2.5 * r ** 4 * (A[3] * B[0] - B[3] * A[0] + A[2] * B[1] - B[2] * A[1]) / (sq3 * 8.0)
    s_xx = (zbs * zas) ** 2.5 * r ** 5 * (A[4] * B[0] + B[4] * A[0] - 2.0 * A[2] * B[2]) / 48.0
    S111 = torch.where(hh, s_hh, torch.where(xh, s_xh, s_xx))
    # p(i)-s(j): zeta_p of hydrogen is 0 -> guard the unused lanes against 0/0
    zap_ = torch.where(hh, torch.ones_like(zap), zap)
    A, B = _AB(r, zap_, zbs)
    t_xh = zbs ** 1.5 * zap_ ** 2.5 * r ** 4 * (A[2] * B[0] - B[2] * A[0] + A[3] * B[1] - B[3] * A[1]) / 8.0
    t_xx = (zbs * zap_) ** 2.5 * r ** 5 * (A[3] * (B[0] - B[2]) - A[1] * (B[2] - B[4])
                                          + B[3] * (A[0] - A[2]) - B[1] * (A[2] - A[4])) / (16.0 * sq3)
    S211 = torch.where(xh, t_xh, torch.where(xx, t_xx, zero))
    zbp_ = torch.where(xx, zbp, torch.ones_like(zbp))
    A, B = _AB(r, zas, zbp_)
    S121 = torch.where(xx, (zbp_ * zas) ** 2.5 * r ** 5 * (A[3] * (B[0] - B[2]) - A[1] * (B[2] - B[4])
                                                           - B[3] * (A[0] - A[2]) + B[1] * (A[2] - A[4])) / (16.0 * sq3), zero)
    zap2 = torch.where(xx, zap, torch.ones_like(zap))
    A, B = _AB(r, zap2, zbp_)
    pre = (zbp_ * zap2) ** 2.5 * r ** 5 / 16.0
    S221 = torch.where(xx, -pre * (B[2] * (A[4] + A[0]) - A[2] * (B[4] + B[0])), zero)
    S222 = torch.where(xx, 0.5 * pre * (A[4] * (B[0] - B[2]) - B[4] * (A[0] - A[2]) - A[2] * B[0] + B[2] * A[0]), zero)
    # direction cosines (:24-45)
    xy = torch.norm(xij[:, :2], dim=1)        # zero sub-gradient on the z axis, as the reference's norm
    gen = xy >= 1.0e-10
    sgn = torch.sign(xij[:, 2]).detach()
    xys = torch.where(gen, xy, torch.ones_like(xy))
    ca = torch.where(gen, xij[:, 0] / xys, sgn)
    sa = torch.where(gen, xij[:, 1] / xys, zero)
    cb = torch.where(gen, xij[:, 2], sgn)
    sb = torch.where(gen, xy, zero)
    u = torch.stack((ca * sb, sa * sb, cb), dim=1)                 # p_mu projection on the axis
    rot = torch.zeros(P, 3, 3, dtype=rij.dtype)                    # (delta - u u^T) written as :207-224
    rot[:, 0, 0] = (ca * cb) ** 2 + sa ** 2
    rot[:, 0, 1] = rot[:, 1, 0] = ca * cb * sa * cb - sa * ca
    rot[:, 0, 2] = rot[:, 2, 0] = -ca * cb * sb
    rot[:, 1, 1] = (sa * cb) ** 2 + ca ** 2
    rot[:, 1, 2] = rot[:, 2, 1] = -sa * cb * sb
    rot[:, 2, 2] = sb ** 2
    di = torch.zeros(P, 4, 4, dtype=rij.dtype)
    di[:, 0, 0] = S111
    di[:, 1:, 0] = S211.unsqueeze(1) * u
    di[:, 0, 1:] = -S121.unsqueeze(1) * u
    di[:, 1:, 1:] = -S221.reshape(-1, 1, 1) * u.unsqueeze(2) * u.unsqueeze(1) + S222.reshape(-1, 1, 1) * rot
    return di


# --- core Hamiltonian (hcore.py:6-167) --------------------------------------
def hcore(ps, par):
    """ps: parse() dict; par: per-atom parameter dict.  Returns M[B*S*S,4,4] (upper only), w."""
    B, S = ps['nmol'], ps['molsize']
    idxi, idxj = ps['idxi'], ps['idxj']
    zeta = torch.stack((par['zeta_s'], par['zeta_p']), dim=1)
    near = ps['rij'] <= OVERLAP_CUTOFF
    di = torch.zeros(idxi.shape[0], 4, 4, dtype=ps['rij'].dtype)
    if near.any():
        di[near] = overlap(ps['ni'][near], ps['nj'][near], ps['xij'][near], ps['rij'][near],
                           zeta[idxi][near], zeta[idxj][near])
    dd, qq, rho0, rho1, rho2 = atom_multipoles(ps['Z'], par['zeta_s'], par['zeta_p'], par['g_ss'],
                                               par['g_pp'], par['g_p2'], par['h_sp'])
    w, e1b, e2a = two_center_integrals(ps['ni'], ps['nj'], ps['xij'], ps['rij'], dd, qq, rho0, rho1, rho2,
                                       idxi, idxj)
    M = torch.zeros(B * S * S, 4, 4, dtype=ps['rij'].dtype)
    diag = torch.zeros(ps['Z'].shape[0], 4, 4, dtype=ps['rij'].dtype)
    diag[:, 0, 0] = par['U_ss']
    diag[:, 1, 1] = diag[:, 2, 2] = diag[:, 3, 3] = par['U_pp']
    diag = diag.index_add(0, idxi, e1b).index_add(0, idxj, e2a)
    M = M.index_put((ps['maskd'],), diag)
    bs, bp = par['beta_s'], par['beta_p']
    bi = torch.stack((bs[idxi], bp[idxi], bp[idxi], bp[idxi]), dim=1)
    bj = torch.stack((bs[idxj], bp[idxj], bp[idxj], bp[idxj]), dim=1)
    M = M.index_put((ps['mask'],), di * (bi.unsqueeze(2) + bj.unsqueeze(1)) / 2.0)
    return M, w


# --- Fock matrix (fock.py:6-139) --------------------------------------------
_WT = torch.tensor([1.0, 2.0, 1.0, 2.0, 2.0, 1.0, 2.0, 2.0, 2.0, 1.0], dtype=torch.float64)
_UR = torch.tensor([n for (m, n) in _PAIRS10])      # upper-triangle rows of the 10 packed products
_UC = torch.tensor([m for (m, n) in _PAIRS10])
_IND_T = torch.tensor(_IND)


def to_blocks(X, B, S):
    return X.reshape(B, S, 4, S, 4).transpose(2, 3).reshape(B * S * S, 4, 4)


def from_blocks(Xb, B, S):
    return Xb.reshape(B, S, S, 4, 4).transpose(2, 3).reshape(B, 4 * S, 4 * S)


def fock(ps, par, P, M, w):
    B, S = ps['nmol'], ps['molsize']
    maskd, mask, idxi, idxj = ps['maskd'], ps['mask'], ps['idxi'], ps['idxj']
    Pb = to_blocks(P, B, S)
    Pd = Pb[maskd]                                  # [A,4,4] diagonal blocks
    gss, gpp, gsp, gp2, hsp = par['g_ss'], par['g_pp'], par['g_sp'], par['g_p2'], par['h_sp']
    pp = Pd[:, 1, 1] + Pd[:, 2, 2] + Pd[:, 3, 3]
    one = torch.zeros_like(Pd)
    one[:, 0, 0] = 0.5 * Pd[:, 0, 0] * gss + pp * (gsp - 0.5 * hsp)
    for i in (1, 2, 3):
        one[:, i, i] = Pd[:, 0, 0] * (gsp - 0.5 * hsp) + 0.5 * Pd[:, i, i] * gpp \
            + (pp - Pd[:, i, i]) * (1.25 * gp2 - 0.25 * gpp)
        one[:, 0, i] = Pd[:, 0, i] * (1.5 * hsp - 0.5 * gsp)
    for i, j in ((1, 2), (1, 3), (2, 3)):
        one[:, i, j] = Pd[:, i, j] * (0.75 * gpp - 1.25 * gp2)
    # two-centre Coulomb: packed, weighted densities of each partner
    pk = Pd[:, _UR, _UC] * _WT                      # [A,10]
    toB = torch.einsum('pk,pkl->pl', pk[idxi], w)   # field of A's density on B's products
    toA = torch.einsum('pl,pkl->pk', pk[idxj], w)
    coul = torch.zeros(Pd.shape[0], 10, dtype=P.dtype).index_add(0, idxi, toA).index_add(0, idxj, toB)
    one[:, _UR, _UC] = one[:, _UR, _UC] + coul
    F = M.index_add(0, maskd, one)
    # exchange on the off-diagonal block
    wx = w[:, _IND_T][:, :, :, _IND_T]              # [P, mu, nu, lam, sig]
    K = torch.einsum('pns,pmnls->pml', -0.5 * Pb[mask], wx)
    F = F.index_add(0, mask, K)
    F0 = from_blocks(F, B, S)
    return F0 + F0.triu(1).transpose(1, 2)


def elec_energy(P, F, Hc):
    """energy.py:19-36; Hc is the upper-only padded core matrix."""
    h = Hc.triu() + Hc.triu(1).transpose(1, 2)
    return 0.5 * (P * (h + F)).sum(dim=(1, 2))


# --- pack / unpack (pack.py:7-41) as index maps -----------------------------
def packed_index(nHeavy, nHydro, S):
    """idx[B, mmax] = padded orbital index of packed orbital q (or -1)."""
    B = nHeavy.shape[0]
    norb = 4 * nHeavy + nHydro
    mmax = int(norb.max())
    q = torch.arange(mmax).unsqueeze(0).expand(B, mmax)
    nh4 = (4 * nHeavy).unsqueeze(1)
    idx = torch.where(q < nh4, q, nh4 + 4 * (q - nh4))
    return torch.where(q < norb.unsqueeze(1), idx, torch.full_like(idx, -1)), norb


def pack(X, nHeavy, nHydro):
    B, n, _ = X.shape
    idx, norb = packed_index(nHeavy, nHydro, n // 4)
    ok = idx >= 0
    g = idx.clamp(min=0)
    Y = X.gather(1, g.unsqueeze(2).expand(B, g.shape[1], n)).gather(2, g.unsqueeze(1).expand(B, g.shape[1], g.shape[1]))
    return Y * (ok.unsqueeze(2) & ok.unsqueeze(1))


def unpack(Y, nHeavy, nHydro, n):
    B, m, _ = Y.shape
    idx, norb = packed_index(nHeavy, nHydro, n // 4)
    ok = idx >= 0
    g = torch.where(ok, idx, torch.full_like(idx, n))          # dump padding into an extra row/col
    X = torch.zeros(B, n + 1, n + 1, dtype=Y.dtype)
    rows = g.unsqueeze(2).expand(B, m, m)
    cols = g.unsqueeze(1).expand(B, m, m)
    X[torch.arange(B).reshape(B, 1, 1).expand(B, m, m), rows, cols] = Y
    return X[:, :n, :n]


# --- density from the Fock matrix -------------------------------------------
def density_eigh(F, nHeavy, nHydro, nocc, want_eig=False):
    """sym_eig_trunc (diag.py:102-172): eigh of the packed matrix (upper triangle), padding rows
    pushed above the spectrum, P = 2 C_occ C_occ^T, unpacked."""
    B, n, _ = F.shape
    x = pack(F, nHeavy, nHydro)
    m = x.shape[1]
    norb = 4 * nHeavy + nHydro
    aii = x.diagonal(dim1=1, dim2=2)
    rad = x.abs().sum(2) - aii.abs()
    hN = (aii + rad).max(1)[0]
    dE = hN - (aii - rad).min(1)[0]
    q = torch.arange(m).unsqueeze(0)
    padq = q >= norb.unsqueeze(1)
    fill = hN.unsqueeze(1) + (1.0 + 0.005 * (q - norb.unsqueeze(1) + 1).to(F.dtype)) * dE.unsqueeze(1)
    x = x + torch.diag_embed(torch.where(padq, fill, torch.zeros_like(fill)))
    e0, v = torch.linalg.eigh(x, UPLO='U')
    occ = (q < nocc.unsqueeze(1)).to(F.dtype)
    t = 2.0 * torch.einsum('bik,bk,bjk->bij', v, occ, v)
    Pn = unpack(t, nHeavy, nHydro, n)
    if want_eig:
        e = torch.zeros(B, n, dtype=F.dtype)
        e[:, :m] = torch.where(padq, torch.zeros_like(e0), e0)
        return Pn, e, v
    return Pn


def orbital_charges(v, nHeavy, nHydro, molsize):
    """`charge` of scf_loop (scf_loop.py:788-800) from the eigenvector array of sym_eig_trunc, literally:
    charge[b, :norb, :nX] = (v^2)[:norb, :4 nX].reshape(norb, 4, nX).sum(1), charge[b, :norb, nX:nX+nH] =
    (v^2)[:norb, 4 nX:4 nX + nH]  (first index of v as stored by eigh, second index the eigenvector)."""
    B = v.shape[0]
    charge = torch.zeros(B, 4 * molsize, molsize, dtype=v.dtype)
    v2 = v ** 2
    for b in range(B):
        nX, nH = int(nHeavy[b]), int(nHydro[b])
        no = 4 * nX + nH
        charge[b, :no, :nX] = v2[b, :no, :4 * nX].reshape(no, 4, nX).sum(dim=1)
        charge[b, :no, nX:nX + nH] = v2[b, :no, 4 * nX:4 * nX + nH]
    return charge


def sp2_density(Fp, nocc, eps):
    """SP2 purification on packed matrices (SP2.py:3-72), fp64 branch.  Returns 2 X."""
    eps = min(max(eps, 1.0e-7), 1.0e-3)
    N, D, _ = Fp.shape
    occ = nocc.to(Fp.dtype)
    aii = Fp.diagonal(dim1=1, dim2=2)
    rad = Fp.abs().sum(2) - aii.abs()
    h1 = (aii - rad).min(1)[0]
    hN = (aii + rad).max(1)[0]
    X = (torch.eye(D, dtype=Fp.dtype).unsqueeze(0) * hN.reshape(-1, 1, 1) - Fp) / (hN - h1).reshape(-1, 1, 1)
    tr = lambda a: a.diagonal(dim1=1, dim2=2).sum(1)
    e0 = (tr(X) - occ).abs()
    e1 = e0.clone()
    live = torch.ones(N, dtype=torch.bool)
    niter = torch.zeros(N, dtype=torch.int64)
    while live.any():
        Xl = X[live]
        X2 = Xl.matmul(Xl)
        t2, t1, o = tr(X2), tr(Xl), occ[live]
        square = (t2 - o).abs() < (2.0 * t1 - t2 - o).abs()
        Xn = torch.where(square.reshape(-1, 1, 1), X2, 2.0 * Xl - X2)
        X[live] = Xn
        e2_prev = e0[live]
        e1[live] = e2_prev
        e0[live] = (tr(Xn) - o).abs()
        niter[live] += 1
        done = (e0[live] < eps) & (e1[live] < eps)
        idx = torch.nonzero(live).reshape(-1)
        live[idx[done]] = False
    return 2.0 * X, niter


# --- SCF drivers (scf_loop.py:32-510, 671-806) ------------------------------
def initial_density(ps):
    """scf_loop.py:700-710."""
    B, S = ps['nmol'], ps['molsize']
    Z = ps['Z']
    d = torch.zeros(Z.shape[0], 4, 4, dtype=torch.float64)
    heavy = Z > 1
    for i in range(4):
        d[heavy, i, i] = TORE[Z[heavy]] / 4.0
    d[Z == 1, 0, 0] = 1.0
    Pb = torch.zeros(B * S * S, 4, 4, dtype=torch.float64)
    Pb[ps['maskd']] = d
    return from_blocks(Pb, B, S)


def scf(ps, par, M, w, P, eps, converger=(2,), sp2=(False,), max_iter=None, trace=None):
    """Returns (P, notconverged, niter[B]).  `converger`: (0, alpha) constant mixing (scf_forward0),
    (1,) adaptive mixing (scf_forward1), (2,) adaptive then Pulay (scf_forward2)."""
    max_iter = MAX_ITER if max_iter is None else max_iter
    B, S = ps['nmol'], ps['molsize']
    n = 4 * S
    nH, nX, nocc = ps['nHydro'], ps['nHeavy'], ps['nocc']
    Hc = from_blocks(M, B, S)
    P = P.clone()

    def density(F, sel):
        if sp2[0]:
            D, _ = sp2_density(pack(F[sel], nX[sel], nH[sel]), nocc[sel], sp2[1])
            return unpack(D, nX[sel], nH[sel], n)
        return density_eigh(F[sel], nX[sel], nH[sel], nocc[sel])

    F = fock(ps, par, P, M, w)
    E = elec_energy(P, F, Hc)
    err = torch.ones(B, dtype=P.dtype)
    nc = torch.ones(B, dtype=torch.bool)
    niter = torch.zeros(B, dtype=torch.int64)
    Pold = torch.zeros_like(P)
    state = dict(k=0)

    def finish_step(F):
        nonlocal E, nc
        En = elec_energy(P[nc], F[nc], Hc[nc])
        err[nc] = (En - E[nc]).abs()
        E[nc] = En
        niter[nc] += 1
        nc = err > eps
        state['k'] += 1
        if trace is not None:
            trace.append((state['k'] - 1, float(err.max()), int(nc.sum())))

    kind = converger[0]
    if kind == 0:
        alpha = converger[1]
        while nc.any():
            P[nc] = alpha * P[nc] + (1.0 - alpha) * density(F, nc)
            F = fock(ps, par, P, M, w)
            finish_step(F)
            if state['k'] >= max_iter:
                break
        return P, nc, niter
    if kind not in (1, 2):
        raise ValueError("scf_converger must be [0, alpha], [1] or [2]")

    def direct():
        nonlocal F
        Pn = density(F, nc)
        Pold[nc] = P[nc]
        P[nc] = Pn
        F = fock(ps, par, P, M, w)
        finish_step(F)

    def adaptive():
        nonlocal F
        Pn = density(F, nc)
        dn = Pn.diagonal(dim1=1, dim2=2)
        dc = P[nc].diagonal(dim1=1, dim2=2)
        do = Pold[nc].diagonal(dim1=1, dim2=2)
        fac = torch.sqrt(((dn - dc) ** 2).sum(1) / ((dn - dc * 2.0 + do) ** 2).sum(1)).reshape(-1, 1, 1)
        Pold[nc] = P[nc]
        P[nc] = (1.0 + fac) * Pn - fac * P[nc]
        F = fock(ps, par, P, M, w)
        finish_step(F)

    for _ in range(2):
        if not nc.any():
            return P, nc, niter
        direct()
    if kind == 1:
        while nc.any():
            adaptive()
            if state['k'] >= max_iter:
                break
        return P, nc, niter
    if not nc.any():
        return P, nc, niter
    adaptive()
    # Pulay / DIIS with a ring of nFock=5 (scf_loop.py:404-510), including the EMAT ring quirk
    nF = 5
    FO = torch.zeros(B, nF, n, n, dtype=P.dtype)
    CM = torch.zeros(B, nF, n, n, dtype=P.dtype)
    EM = (torch.eye(nF + 1, dtype=P.dtype) - 1.0).tril().unsqueeze(0).repeat(B, 1, 1)
    counter, cF = -1, 0

    def store():
        nonlocal counter, cF
        cF = min(cF + 1, nF)
        counter = (counter + 1) % nF
        FO[nc, counter] = F[nc]
        CM[nc, counter] = (F[nc].matmul(P[nc]) - P[nc].matmul(F[nc])).triu()
        EM[nc, counter, :cF] = (CM[nc, counter:counter + 1] * CM[nc, :cF]).sum(dim=(2, 3))

    while cF < 2:
        if not nc.any():
            return P, nc, niter
        store()
        P[nc] = density(F, nc)
        F = fock(ps, par, P, M, w)
        finish_step(F)
    while nc.any():
        EV_ = EM[nc] + EM[nc].tril(-1).transpose(1, 2)
        EV_[:, :cF, :cF] = EV_[:, :cF, :cF] / EV_[:, counter:counter + 1, counter:counter + 1]
        coeff = -torch.inverse(EV_[:, :cF + 1, :cF + 1])[:, :-1, -1]
        F[nc] = (FO[nc, :cF] * coeff.reshape(-1, cF, 1, 1)).sum(1)
        P[nc] = density(F, nc)
        F = fock(ps, par, P, M, w)
        store()
        finish_step(F)
        if state['k'] >= max_iter:
            break
    return P, nc, niter


# --- energies (energy.py:4-118) ---------------------------------------------
def pair_nuclear_energy(ps, tab_atom, method, gam):
    ni, nj, idxi, idxj = ps['ni'], ps['nj'], ps['idxi'], ps['idxj']
    R = ps['rij'] * A0
    alpha = tab_atom['alpha']
    zz = TORE[ni] * TORE[nj]
    ea = torch.exp(-alpha[idxi] * R)
    nhoh = ((ni == 7) | (ni == 8)) & (nj == 1)
    ea = torch.where(nhoh, ea * R, ea)
    eb = torch.exp(-alpha[idxj] * R)
    E = zz * gam * (1.0 + ea + eb)
    if method in ('AM1', 'PM3'):
        K, L, Mg = tab_atom['K'], tab_atom['L'], tab_atom['M']
        Rc = R.reshape(-1, 1)
        ga = (K[idxi] * torch.exp(-L[idxi] * (Rc - Mg[idxi]) ** 2)).sum(1)
        gb = (K[idxj] * torch.exp(-L[idxj] * (Rc - Mg[idxj]) ** 2)).sum(1)
        E = E + zz / R * (ga + gb)
    elif method != 'MNDO':
        raise ValueError("Supported Method: MNDO, AM1, PM3")
    return E


def isolated_atom_energy(Z, par):
    return par['U_ss'] * USSC[Z] + par['U_pp'] * UPPC[Z] + par['g_ss'] * GSSC[Z] + par['g_pp'] * GPPC[Z] \
        + par['g_sp'] * GSPC[Z] + par['g_p2'] * GP2C[Z] + par['h_sp'] * HSPC[Z]


def run(species, coordinates, method='AM1', scf_eps=1.0e-6, converger=(2,), sp2=(False,), cutoff=1.0e10,
        P0=None, forces=True, max_iter=None, Hf_flag=True, trace=None):
    """One `Force`/`Energy` call (basics.py:271-390).  Returns a dict of reference-named results."""
    coords = coordinates.detach().clone().requires_grad_(forces)
    ps = parse(species, coords, cutoff)
    tab = load_params(method)
    par = {k: v[ps['Z']] for k, v in tab.items()}
    M, w = hcore(ps, par)
    P = initial_density(ps) if P0 is None else P0
    with torch.no_grad():
        P, nc, niter = scf(ps, par, M.detach(), w.detach(), P, scf_eps, converger, sp2, max_iter, trace)
    F = fock(ps, par, P, M, w)
    Hc = from_blocks(M, ps['nmol'], ps['molsize'])
    Eelec = elec_energy(P, F, Hc)
    EnucAB = pair_nuclear_energy(ps, par, method, w[:, 0, 0])
    B = ps['nmol']
    Enuc = torch.zeros(B, dtype=torch.float64).index_add(0, ps['pair_molid'], EnucAB)
    Etot = Eelec + Enuc
    Eiso = torch.zeros(B, dtype=torch.float64).index_add(0, ps['atom_molid'], isolated_atom_energy(ps['Z'], par))
    Hf = Etot - Eiso
    if Hf_flag:
        Hf = Hf + torch.zeros(B, dtype=torch.float64).index_add(0, ps['atom_molid'], EHEAT[ps['Z']])
    out = dict(ps=ps, M=M.detach(), w=w.detach(), P=P, F=F.detach(), Hcore=Hc.detach(), Eelec=Eelec.detach(),
               EnucAB=EnucAB.detach(), Enuc=Enuc.detach(), Etot=Etot.detach(), Eiso=Eiso, Hf=Hf.detach(),
               notconverged=nc, niter=niter)
    if forces:
        Hf.sum().backward()
        out['force'] = -coords.grad.detach()
    return out
