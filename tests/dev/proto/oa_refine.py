"""CPU prototype: Ogita-Aishima iterative refinement (SIAM J. Matrix Anal. 2018) of the previous SCF iteration's
eigenvectors as the density solver inside the oracle's SCF loop (one molecule at a time).  Reports refinement steps
per call, fallbacks, and the density / energy / iteration-count deltas against the eigh path."""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'oracle'))
import seqm_oracle as orc
from pyseqm_b200 import synth
torch.set_default_dtype(torch.float64)

STORE = {}
STATS = {}
REF_EIGH = orc.density_eigh
TOL = float(os.environ.get('OA_TOL', '1e-7'))


THETA = float(os.environ.get('OA_THETA', '0.1'))


def oa_refine(A, X, nocc, max_steps=8, tol=TOL):
    m = A.shape[0]
    prev = None
    for step in range(1, max_steps + 1):
        T = A @ X
        S = X.T @ T
        R = np.eye(m) - X.T @ X
        lam = np.diag(S) / (1.0 - np.diag(R))
        dl = lam[None, :] - lam[:, None]
        num = S + lam[None, :] * R
        skip = np.abs(num) > THETA * np.abs(dl)          # rotation angle not small: leave the pair alone
        with np.errstate(divide='ignore', invalid='ignore'):
            E = num / dl
        E = np.where(skip, R / 2.0, E)
        np.fill_diagonal(E, np.diag(R) / 2.0)
        order = np.argsort(lam, kind='stable')
        occ = np.zeros(m, bool); occ[order[:nocc]] = True
        ov = occ[:, None] != occ[None, :]
        np.fill_diagonal(skip, False)
        if (skip & ov).any():
            return None, 'straddle', step
        en = np.abs(E).max()
        if step == 1:
            STATS['e0'].append(en)
        if prev is not None and en > prev:
            return None, 'diverge', step
        X = X + X @ E
        if en < tol:
            return (X, lam, occ), 'ok', step
        prev = en
    return None, 'maxsteps', max_steps


def density_oa(F, nHeavy, nHydro, nocc, want_eig=False):
    B, n, _ = F.shape
    assert B == 1
    x = orc.pack(F, nHeavy, nHydro)
    Pref = REF_EIGH(F, nHeavy, nHydro, nocc)
    m = int(4 * nHeavy[0] + nHydro[0])
    A = x[0, :m, :m].numpy()
    A = np.triu(A) + np.triu(A, 1).T
    no = int(nocc[0])
    STATS['calls'] += 1
    res = None
    if 0 in STORE:
        res, why, steps = oa_refine(A, STORE[0], no)
        if res is None:
            STATS['why'].append((why, steps, STATS['k']))
        else:
            STATS['steps'].append(steps)
    if res is None:
        lam, X = np.linalg.eigh(A)
        occ = np.zeros(m, bool); occ[:no] = True
    else:
        X, lam, occ = res
    STATS['k'] += 1
    STORE[0] = X
    C = X[:, occ]
    out = torch.zeros_like(x)
    out[0, :m, :m] = torch.as_tensor(2.0 * C @ C.T)
    Pn = orc.unpack(out, nHeavy, nHydro, n)
    STATS['maxerr'] = max(STATS['maxerr'], float((Pn - Pref).abs().max()))
    return Pn


def study(name, sp, xyz, method):
    tot = dict(calls=0, steps=[], why=[], maxerr=0.0, dE=0.0, dit=0, e0=[])
    for b in range(sp.shape[0]):
        species, coords = torch.as_tensor(sp[b:b + 1]), torch.as_tensor(xyz[b:b + 1])
        orc.density_eigh = REF_EIGH
        o0 = orc.run(species, coords, method=method, scf_eps=1e-6, converger=(2, 0.0), forces=False)
        STORE.clear(); STATS.clear(); STATS.update(calls=0, steps=[], why=[], maxerr=0.0, e0=[], k=0)
        orc.density_eigh = density_oa
        o1 = orc.run(species, coords, method=method, scf_eps=1e-6, converger=(2, 0.0), forces=False)
        orc.density_eigh = REF_EIGH
        tot['calls'] += STATS['calls']; tot['steps'] += STATS['steps']; tot['why'] += STATS['why']; tot['e0'] += STATS['e0']
        tot['maxerr'] = max(tot['maxerr'], STATS['maxerr'])
        tot['dE'] = max(tot['dE'], float((o0['Etot'] - o1['Etot']).abs().max()))
        tot['dit'] = max(tot['dit'], int((o0['niter'] - o1['niter']).abs().max()))
    st = np.array(tot['steps'])
    print(name, 'calls', tot['calls'], 'refined', len(st), 'mean steps %.2f' % st.mean(), 'hist', np.bincount(st),
          'fallbacks', len(tot['why']), 'max|dP| %.2e' % tot['maxerr'], 'max|dE| %.2e' % tot['dE'], 'max dit', tot['dit'])
    from collections import Counter
    print('   fallback reasons', Counter((w, k) for w, s, k in tot['why']).most_common(12))
    print('   first-step |E| quantiles', np.quantile(np.array(tot['e0']), [0.1, 0.5, 0.9, 0.99]))


if __name__ == '__main__':
    sp, xyz = synth.benzene(24, seed=0)
    study('benzene AM1', sp, xyz, 'AM1')
    sp, xyz = synth.mixed_organics_sweep(48, seed=0, base=48)
    study('organics MNDO', sp, xyz, 'MNDO')
