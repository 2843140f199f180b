"""Build-container script: run the UNMODIFIED reference (/root/reference) on the LARGE parity cases of
BASELINE.json configs 3-5 and write compact fixtures under tests/golden_big/.

    python tools/make_golden_big.py [name-substring ...]

The full P and F matrices of these cases would be tens of MB, so the fixtures hold per-molecule results (Etot,
Eelec, Hf, Enuc, Eiso, forces, notconverged, the SCF trace) and, for the matrices, the diagonal of P plus the
projections P r and F r on the fixed vector r_i = 1 + 0.5 sin(i + 1) (a linear functional of every row: a 1e-9
error anywhere in P or F shows up at that size).  Cases marked `eig` also store the orbital energies and the
per-orbital `charge` array of scf_loop.py:779-804.
"""
import contextlib
import io
import os
import sys
import time
import warnings

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
from ref_loader import load_reference  # noqa: E402
from pyseqm_b200 import synth  # noqa: E402

torch.set_default_dtype(torch.float64)
warnings.filterwarnings('ignore')
OUT = os.path.join(ROOT, 'tests', 'golden_big')
ELEMENTS = [0, 1, 6, 7, 8]


def probe_vector(n):
    return 1.0 + 0.5 * np.sin(np.arange(n, dtype=np.float64) + 1.0)


def cases():
    out = []
    sp, xyz = synth.mixed_organics_sweep(256, seed=11, base=256)      # config 4's recipe: 5-30 atoms, m = 8..56
    out.append(dict(name='organics256_MNDO_eigh_c2', species=sp, coords=xyz, method='MNDO', eps=1e-6, conv=[2, 0.0],
                    sp2=[False]))
    sp, xyz = synth.mixed_organics(32, seed=12, min_atoms=31, max_atoms=60, attempts=5000)      # m = 55..106
    out.append(dict(name='organics32big_MNDO_eigh_c2', species=sp, coords=xyz, method='MNDO', eps=1e-6,
                    conv=[2, 0.0], sp2=[False]))
    sp, xyz = synth.mixed_organics(6, seed=13, min_atoms=62, max_atoms=84, attempts=20000)      # m above 118
    out.append(dict(name='organics6huge_MNDO_eigh_c2', species=sp, coords=xyz, method='MNDO', eps=1e-6,
                    conv=[2, 0.0], sp2=[False]))
    sp, xyz = synth.c60(4, seed=5)
    out.append(dict(name='c60x4_PM3_sp2_c2', species=sp, coords=xyz, method='PM3', eps=1e-6, conv=[2, 0.0],
                    sp2=[True, 1e-5]))
    sp, xyz = synth.c60(2, seed=6)
    out.append(dict(name='c60x2_PM3_eigh_c2', species=sp, coords=xyz, method='PM3', eps=1e-6, conv=[2, 0.0],
                    sp2=[False]))
    sp, xyz = synth.water_cluster(120, seed=7)
    out.append(dict(name='water120_PM3_sp2_c2', species=sp, coords=xyz, method='PM3', eps=1e-6, conv=[2, 0.0],
                    sp2=[True, 1e-5]))
    sp, xyz = synth.small_molecules()
    out.append(dict(name='small_MNDO_eig', species=sp, coords=xyz, method='MNDO', eps=1e-8, conv=[2, 0.0],
                    sp2=[False], eig=True))
    sp, xyz = synth.benzene(4, seed=21)
    out.append(dict(name='benzene4_AM1_eig', species=sp, coords=xyz, method='AM1', eps=1e-8, conv=[2, 0.0],
                    sp2=[False], eig=True))
    sp, xyz = synth.mixed_organics(8, seed=3, min_atoms=5, max_atoms=14)
    out.append(dict(name='mixed8_PM3_eig', species=sp, coords=xyz, method='PM3', eps=1e-8, conv=[2, 0.0],
                    sp2=[False], eig=True))
    sp, xyz = synth.mixed_organics(8, seed=3, min_atoms=5, max_atoms=14)
    out.append(dict(name='mixed8_PM3_sp2', species=sp, coords=xyz, method='PM3', eps=1e-6, conv=[2, 0.0],
                    sp2=[True, 1e-5]))
    sp, xyz = synth.water_cluster(500, seed=0)      # BASELINE.json configs[4]; the reference needs ~10 min of CPU (8 cores)
    out.append(dict(name='water500_PM3_sp2_c2', species=sp, coords=xyz, method='PM3', eps=1e-6, conv=[2, 0.0],
                    sp2=[True, 1e-5], slow=True))
    return out


def run_reference(seqm, case):
    from seqm.basics import Force
    from seqm.seqm_functions.constants import Constants
    from seqm.seqm_functions import scf_loop as ref_scf
    species = torch.as_tensor(case['species'])
    coords = torch.as_tensor(case['coords']).clone()
    const = Constants()
    sp = dict(method=case['method'], scf_eps=case['eps'], scf_converger=case['conv'], sp2=case['sp2'],
              elements=ELEMENTS, learned=[], parameter_file_dir=os.path.join(seqm.__path__[0], 'params') + '/',
              pair_outer_cutoff=1.0e10, eig=bool(case.get('eig', False)))
    ref_scf.debug = True
    buf = io.StringIO()
    t0 = time.time()
    fm = Force(sp)
    with contextlib.redirect_stdout(buf):
        force, P, Etot, Hf, Eelec, Enuc, Eiso, EnucAB, e, charge, notconv = fm(const, coords, species)
    dt = time.time() - t0
    ref_scf.debug = False
    trace = []
    for line in buf.getvalue().splitlines():
        t = line.split()
        if len(t) == 4 and t[0] == 'scf':
            trace.append((int(t[1]), float(t[2]), int(t[3])))
    # final Fock matrix at the converged density (what Hamiltonian returns as F)
    from seqm.basics import Parser, Pack_Parameters
    from seqm.seqm_functions.hcore import hcore
    from seqm.seqm_functions.fock import fock
    with torch.no_grad():
        ps = Parser(sp)(const, species, coords.detach())
        nmol, molsize, nHeavy, nHydro, nocc, Z, maskd, atom_molid, mask, pair_molid, ni, nj, idxi, idxj, xij, rij = ps
        par = Pack_Parameters(sp)(Z, learned_params=dict())
        beta = torch.stack((par['beta_s'], par['beta_p']), dim=1)
        M, w = hcore(const, nmol, molsize, maskd, mask, idxi, idxj, ni, nj, xij, rij, Z, par['zeta_s'],
                     par['zeta_p'], par['U_ss'], par['U_pp'], par['g_ss'], par['g_pp'], par['g_p2'], par['h_sp'],
                     beta)
        F = fock(nmol, molsize, P.detach(), M, maskd, mask, idxi, idxj, w, par['g_ss'], par['g_pp'], par['g_sp'],
                 par['g_p2'], par['h_sp'])
    n = P.shape[-1]
    r = torch.as_tensor(probe_vector(n))
    Pd = P.detach()
    g = dict(species=case['species'], coords=case['coords'], method=case['method'], eps=case['eps'],
             conv=np.array(case['conv'], dtype=np.float64), sp2=np.array([float(x) for x in case['sp2']]),
             eig=np.array(bool(case.get('eig', False))),
             force=force.numpy(), Etot=Etot.detach().numpy(), Hf=Hf.detach().numpy(), Eelec=Eelec.detach().numpy(),
             Enuc=Enuc.detach().numpy(), Eiso=Eiso.detach().numpy(), notconverged=notconv.numpy(),
             trace=np.array(trace, dtype=np.float64), ref_seconds=dt,
             Pdiag=Pd.diagonal(dim1=1, dim2=2).numpy().copy(), Pr=(Pd @ r).numpy(), Fr=(F @ r).numpy(),
             Ffro=F.pow(2).sum(dim=(1, 2)).sqrt().numpy(), w_sum=float(w.sum()), nHeavy=nHeavy.numpy(),
             nHydro=nHydro.numpy(), nocc=nocc.numpy())
    if case.get('eig'):
        g['e'] = e.detach().numpy()
        g['charge'] = charge.detach().numpy()
    return g


def main():
    os.makedirs(OUT, exist_ok=True)
    seqm = load_reference()
    only = sys.argv[1:]
    for case in cases():
        if only and not any(s in case['name'] for s in only):
            continue
        if case.get('slow') and not only:
            continue
        t0 = time.time()
        g = run_reference(seqm, case)
        np.savez_compressed(os.path.join(OUT, case['name'] + '.npz'), **g)
        print(case['name'], 'ref %.1fs (total %.1fs)' % (g['ref_seconds'], time.time() - t0), 'iters', len(g['trace']),
              'notconverged', int(g['notconverged'].sum()), 'Etot[0]', float(g['Etot'][0]), flush=True)


if __name__ == '__main__':
    main()
