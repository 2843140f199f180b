"""CPU emulation of the warp kernel's cyclic Jacobi (odd-even transposition ordering, warm start from the previous SCF
iteration's eigenvectors, stop rule: one last sweep once off^2 <= tol tot^2): sweeps per SCF iteration."""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'oracle'))
import seqm_oracle as orc
from pyseqm_b200 import synth
torch.set_default_dtype(torch.float64)
REF = orc.density_eigh
TOL = float(os.environ.get('JTOL', '1e-18'))
LOG = []
STORE = {}


def jacobi(A, V, tol):
    n = A.shape[0]
    N = n + (n & 1)
    Ap = np.zeros((N, N)); Ap[:n, :n] = A
    Vp = np.eye(N); Vp[:n, :n] = V
    A = Vp.T @ Ap @ Vp
    V = Vp
    pos = list(range(N))
    sweeps, last = 0, False
    while sweeps < 40:
        off = (A ** 2).sum() - (np.diag(A) ** 2).sum(); tot = (A ** 2).sum()
        if off == 0 or last: break
        if off <= tol * tot: last = True
        for r in range(N):
            start = r & 1
            for k in range(start, N - 1, 2):
                p, q = pos[k], pos[k + 1]
                if A[p, q] != 0.0:
                    d = A[q, q] - A[p, p]; b = 2 * A[p, q]
                    t = (np.sign(d) if d != 0 else 1.0) * b / (abs(d) + np.hypot(d, b))
                    c = 1 / np.sqrt(1 + t * t); s = t * c
                    J = np.eye(N); J[p, p] = c; J[q, q] = c; J[p, q] = s; J[q, p] = -s
                    A = J.T @ A @ J; V = V @ J
                pos[k], pos[k + 1] = q, p
        sweeps += 1
    return np.diag(A)[:n], V[:n, :n], sweeps


def dens(F, nHeavy, nHydro, nocc, want_eig=False):
    x = orc.pack(F, nHeavy, nHydro)
    m = int(4 * nHeavy[0] + nHydro[0])
    A = x[0, :m, :m].numpy(); A = np.triu(A) + np.triu(A, 1).T
    V0 = STORE.get(0, np.eye(m))
    lam, V, sw = jacobi(A, V0, TOL)
    LOG[-1].append(sw)
    STORE[0] = V
    order = np.argsort(lam, kind='stable')[:int(nocc[0])]
    C = V[:, order]
    out = torch.zeros_like(x); out[0, :m, :m] = torch.as_tensor(2 * C @ C.T)
    return orc.unpack(out, nHeavy, nHydro, F.shape[1])


def study(name, sp, xyz, method):
    for b in range(sp.shape[0]):
        STORE.clear(); LOG.append([])
        orc.density_eigh = dens
        orc.run(torch.as_tensor(sp[b:b + 1]), torch.as_tensor(xyz[b:b + 1]), method=method, scf_eps=1e-6, converger=(2, 0.0), forces=False)
        orc.density_eigh = REF
    L = max(len(l) for l in LOG)
    per = [np.mean([l[k] for l in LOG if len(l) > k]) for k in range(L)]
    print(name, 'sweeps per SCF iteration:', ' '.join('%.1f' % v for v in per), '| total/mol %.1f' % np.mean([sum(l) for l in LOG]),
          'calls/mol %.1f' % np.mean([len(l) for l in LOG]))
    LOG.clear()


if __name__ == '__main__':
    sp, xyz = synth.benzene(6, seed=0)
    study('benzene', sp, xyz, 'AM1')
    sp, xyz = synth.mixed_organics_sweep(8, seed=0, base=8)
    study('organics', sp, xyz, 'MNDO')
