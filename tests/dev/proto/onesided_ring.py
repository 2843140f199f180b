"""Prototype (numpy, CPU) of the one-sided Jacobi with the recursive ring ordering used by the warp eigensolver:
G = (A + sigma I) V0, columns orthogonalised pair by pair; K processors hold columns (P_k, Q_k); phase of group size
Kg: Kg rounds, Q shifts by one processor inside the group between rounds; then the two halves of a group exchange
(P's to the first half, Q's to the second).  Counts sweeps and checks the density against eigh on the Fock matrices
of an oracle SCF run.  Test infrastructure only (uses oracle/)."""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'oracle'))
torch.set_default_dtype(torch.float64)
import seqm_oracle as orc
from pyseqm_b200 import synth


THR = 1.0e-13
TMIN = 1.0e-3


def schedule(K):
    """list of rounds: each round = list of (colP, colQ) by slot bookkeeping -> returns ops: ('rot',), ('shift',Kg), ('xchg',Kg)"""
    ops = []
    Kg = K
    while Kg >= 1:
        for r in range(Kg):
            ops.append(('rot',))
            if r + 1 < Kg:
                ops.append(('shift', Kg))
        if Kg > 1:
            ops.append(('xchg', Kg))
        Kg //= 2
    return ops


def jacobi_onesided(A, V0, K, tol=1e-18, fp32_tangent=True, maxsweeps=40):
    m = A.shape[0]
    aii = np.diag(A); rad = np.abs(A).sum(1) - np.abs(aii)
    gl, gu = (aii - rad).min(), (aii + rad).max()
    w = gu - gl
    sigma = -gl + w
    ncol = 2 * K
    G = np.zeros((m, ncol))
    if V0 is None:
        G[:, :m] = A + sigma * np.eye(m)
    else:
        G[:, :V0.shape[1]] = (A + sigma * np.eye(m)) @ V0
    Pslot = list(range(0, ncol, 2)); Qslot = list(range(1, ncol, 2))   # column index held by processor k
    ops = schedule(K)
    sweeps = 0
    nrot_total = 0
    while sweeps < maxsweeps:
        alpha = (G * G).sum(0)
        tot2 = (alpha ** 2).sum()
        off2 = 0.0
        for op in ops:
            if op[0] == 'rot':
                for k in range(K):
                    p, q = Pslot[k], Qslot[k]
                    gam = G[:, p] @ G[:, q]
                    off2 += 2 * gam * gam
                    a, b = alpha[p], alpha[q]
                    small = gam * gam <= THR * THR * a * b   # columns orthogonal to rounding AND a large angle: no rotation (a
                    if gam == 0.0:                           # noise angle inside a degenerate cluster re-mixes its couplings)
                        continue
                    if fp32_tangent:
                        df = np.float32(b - a); bf = np.float32(2 * gam)
                        h = np.sqrt(df * df + bf * bf)
                        den = np.abs(df) + h
                        t = float((bf if df >= 0 else -bf) / den) if den > 0 else 0.0
                    else:
                        z = (b - a) / (2 * gam)
                        t = np.sign(z) / (abs(z) + np.sqrt(1 + z * z)) if z != 0 else 1.0
                    if small and abs(t) > TMIN:
                        continue
                    c = 1 / np.sqrt(1 + t * t); s = c * t
                    gp, gq = G[:, p].copy(), G[:, q].copy()
                    G[:, p] = c * gp - s * gq
                    G[:, q] = s * gp + c * gq
                    # valid for an inexact (FP32) tangent as well: the exact-t shortcut a - t gam loses the difference of
                    # (nearly) degenerate columns
                    alpha[p] = c * c * (a - 2 * t * gam + t * t * b)
                    alpha[q] = c * c * (b + 2 * t * gam + t * t * a)
                    nrot_total += 1
            elif op[0] == 'shift':
                Kg = op[1]
                newQ = list(Qslot)
                for k in range(K):
                    base = k - (k % Kg)
                    newQ[k] = Qslot[base + ((k - base - 1) % Kg)]
                Qslot = newQ
            else:
                Kg = op[1]; h = Kg // 2
                for k in range(K):
                    if (k % Kg) < h:
                        kp = k + h
                        Qslot[k], Pslot[kp] = Pslot[kp], Qslot[k]
        sweeps += 1
        if off2 <= tol * tot2:
            break
    alpha = (G * G).sum(0)
    d = np.sqrt(alpha)
    real = alpha > 0
    V = np.zeros_like(G); V[:, real] = G[:, real] / d[real]
    lam = d - sigma
    return lam, V, real, sweeps


def density(lam, V, real, nocc):
    idx = np.where(real)[0]
    order = idx[np.argsort(lam[idx], kind='stable')]
    C = V[:, order[:nocc]]
    return 2 * C @ C.T


def main():
    nmol = int(sys.argv[1]) if len(sys.argv) > 1 else 3
    sp_, xyz = synth.benzene(nmol, seed=7)
    rec = []
    orig = orc.density_eigh
    def spy(F, nHeavy, nHydro, nocc, want_eig=False):
        rec.append((orc.pack(F, nHeavy, nHydro).numpy().copy(), nocc.numpy().copy()))
        return orig(F, nHeavy, nHydro, nocc, want_eig)
    orc.density_eigh = spy
    orc.run(torch.as_tensor(sp_), torch.as_tensor(xyz), method='AM1', scf_eps=1e-6, converger=(2, 0.0))
    orc.density_eigh = orig
    print("eigensolves recorded:", len(rec), "batch shapes", [r[0].shape for r in rec][:3])
    for fp32 in (True, False):
        tot_sw = 0; worst = 0.0
        for b in range(nmol):
            V0 = None; sw = []
            for (Fp, nocc) in rec:
                if b >= Fp.shape[0]:
                    continue
                A = Fp[b]; m = A.shape[0]
                lam, V, real, s = jacobi_onesided(A, V0, 16, fp32_tangent=fp32)
                P = density(lam, V, real, int(nocc[b]))
                e, U = np.linalg.eigh(A)
                Pref = 2 * U[:, :int(nocc[b])] @ U[:, :int(nocc[b])].T
                worst = max(worst, np.abs(P - Pref).max(), np.abs(np.sort(lam[real]) - e).max())
                V0 = V
                sw.append(s)
            print("fp32 tangent" if fp32 else "fp64 tangent", "mol", b, "sweeps per eigensolve", sw, "sum", sum(sw))
            tot_sw += sum(sw)
        print("  total sweeps", tot_sw, " worst |dP| or |de|: %.2e" % worst)


if __name__ == '__main__':
    main()


def tol_scan():
    sp_, xyz = synth.benzene(4, seed=11)
    rec = []
    orig = orc.density_eigh
    def spy(F, nHeavy, nHydro, nocc, want_eig=False):
        rec.append((orc.pack(F, nHeavy, nHydro).numpy().copy(), nocc.numpy().copy()))
        return orig(F, nHeavy, nHydro, nocc, want_eig)
    orc.density_eigh = spy
    orc.run(torch.as_tensor(sp_), torch.as_tensor(xyz), method='AM1', scf_eps=1e-6, converger=(2, 0.0))
    orc.density_eigh = orig
    for tol in (1e-18, 1e-16, 1e-14, 1e-12):
        tot_sw = 0; worst = 0.0
        for b in range(4):
            V0 = None
            for (Fp, nocc) in rec:
                A = Fp[b]
                lam, V, real, s = jacobi_onesided(A, V0, 16, tol=tol)
                P = density(lam, V, real, int(nocc[b]))
                e, U = np.linalg.eigh(A)
                Pref = 2 * U[:, :int(nocc[b])] @ U[:, :int(nocc[b])].T
                worst = max(worst, np.abs(P - Pref).max())
                V0 = V; tot_sw += s
        print("tol %.0e: sweeps per molecule %.1f  worst |dP| %.2e" % (tol, tot_sw / 4, worst))
