"""Build-container script: run the UNMODIFIED reference (/root/reference) on the parity cases, write the
golden fixtures under tests/golden/ and report oracle-vs-reference deltas.

    python tools/make_golden.py            # regenerate every fixture
The fixtures are what pins the oracle (tests/test_oracle_golden.py) and the CUDA path (tests -m gpu) on the
GPU box, where /root/reference does not exist.
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
OUT = os.path.join(ROOT, 'tests', 'golden')

ELEMENTS = [0, 1, 6, 7, 8]


def cases():
    sp_f, xyz_f = synth.formaldehyde_pair()
    sp_b, xyz_b = synth.benzene(8, seed=1)
    sp_s, xyz_s = synth.small_molecules()
    sp_m, xyz_m = synth.mixed_organics(12, seed=3, min_atoms=5, max_atoms=14)
    sp_w, xyz_w = synth.water_cluster(6, seed=2)
    sp_c, xyz_c = synth.c60(1, seed=4)
    out = []
    for method in ('MNDO', 'AM1', 'PM3'):
        for conv, ctag in (([2, 0.0], 'c2'), ([1], 'c1'), ([0, 0.2], 'c0')):
            if conv[0] != 2 and method != 'MNDO':
                continue
            out.append(dict(name='h2co_%s_eigh_%s' % (method, ctag), species=sp_f, coords=xyz_f, method=method,
                            eps=1e-8, conv=conv, sp2=[False], keep_w=True))
        out.append(dict(name='h2co_%s_sp2_c2' % method, species=sp_f[:1], coords=xyz_f[:1], method=method,
                        eps=1e-8, conv=[2, 0.0], sp2=[True, 1e-7], keep_w=False))
    out.append(dict(name='benzene8_AM1_eigh_c2', species=sp_b, coords=xyz_b, method='AM1', eps=1e-6,
                    conv=[2, 0.0], sp2=[False], keep_w=True))
    out.append(dict(name='benzene8_AM1_sp2_c2', species=sp_b, coords=xyz_b, method='AM1', eps=1e-6,
                    conv=[2, 0.0], sp2=[True, 1e-5], keep_w=False))
    for method in ('MNDO', 'PM3'):
        out.append(dict(name='small_%s_eigh_c2' % method, species=sp_s, coords=xyz_s, method=method, eps=1e-8,
                        conv=[2, 0.0], sp2=[False], keep_w=True))
    out.append(dict(name='small_PM3_eigh_c1', species=sp_s, coords=xyz_s, method='PM3', eps=1e-10,
                    conv=[1], sp2=[False], keep_w=False))
    out.append(dict(name='mixed12_MNDO_eigh_c2', species=sp_m, coords=xyz_m, method='MNDO', eps=1e-6,
                    conv=[2, 0.0], sp2=[False], keep_w=False))
    out.append(dict(name='water6_PM3_sp2_c2', species=sp_w, coords=xyz_w, method='PM3', eps=1e-6,
                    conv=[2, 0.0], sp2=[True, 1e-5], keep_w=False))
    out.append(dict(name='water6_PM3_eigh_c2', species=sp_w, coords=xyz_w, method='PM3', eps=1e-6,
                    conv=[2, 0.0], sp2=[False], keep_w=False))
    out.append(dict(name='c60_PM3_sp2_c2', species=sp_c, coords=xyz_c, method='PM3', eps=1e-6,
                    conv=[2, 0.0], sp2=[True, 1e-5], keep_w=False))
    return out


def run_reference(seqm, case):
    from seqm.basics import Force, Parser
    from seqm.seqm_functions.constants import Constants
    from seqm.seqm_functions import scf_loop as ref_scf
    species = torch.as_tensor(case['species'])
    coords = torch.as_tensor(case['coords']).clone()
    const = Constants()
    sp = dict(method=case['method'], scf_eps=case['eps'], scf_converger=case['conv'], sp2=case['sp2'],
              elements=ELEMENTS, learned=[], parameter_file_dir=os.path.join(seqm.__path__[0], 'params') + '/',
              pair_outer_cutoff=1.0e10)
    ref_scf.debug = True
    buf = io.StringIO()
    t0 = time.time()
    with contextlib.redirect_stdout(buf):
        force, P, Etot, Hf, Eelec, Enuc, Eiso, EnucAB, e, charge, notconv = Force(sp)(const, coords, species)
    dt = time.time() - t0
    ref_scf.debug = False
    trace = []
    for line in buf.getvalue().splitlines():
        t = line.split()
        if len(t) == 4 and t[0] == 'scf':
            trace.append((int(t[1]), float(t[2]), int(t[3])))
    with torch.no_grad():
        ps = Parser(sp)(const, species, coords.detach())
    # intermediates: hcore and the final Fock matrix
    from seqm.basics import Pack_Parameters
    from seqm.seqm_functions.hcore import hcore
    from seqm.seqm_functions.fock import fock
    nmol, molsize, nHeavy, nHydro, nocc, Z, maskd, atom_molid, mask, pair_molid, ni, nj, idxi, idxj, xij, rij = ps
    par = Pack_Parameters(sp)(Z, learned_params=dict())
    with torch.no_grad():
        beta = torch.stack((par['beta_s'], par['beta_p']), dim=1)
        M, w = hcore(const, nmol, molsize, maskd, mask, idxi, idxj, ni, nj, xij, rij, Z, par['zeta_s'],
                     par['zeta_p'], par['U_ss'], par['U_pp'], par['g_ss'], par['g_pp'], par['g_p2'], par['h_sp'],
                     beta)
        F = fock(nmol, molsize, P.detach(), M, maskd, mask, idxi, idxj, w, par['g_ss'], par['g_pp'], par['g_sp'],
                 par['g_p2'], par['h_sp'])
    g = dict(species=case['species'], coords=case['coords'], method=case['method'], eps=case['eps'],
             conv=np.array(case['conv'], dtype=np.float64), sp2=np.array([float(x) for x in case['sp2']]),
             nHeavy=nHeavy.numpy(), nHydro=nHydro.numpy(), nocc=nocc.numpy(), Z=Z.numpy(), maskd=maskd.numpy(),
             atom_molid=atom_molid.numpy(), mask=mask.numpy(), pair_molid=pair_molid.numpy(), ni=ni.numpy(),
             nj=nj.numpy(), idxi=idxi.numpy(), idxj=idxj.numpy(), xij=xij.numpy(), rij=rij.numpy(),
             M=M.numpy(), F=F.numpy(), P=P.detach().numpy(), force=force.numpy(), Etot=Etot.detach().numpy(),
             Hf=Hf.detach().numpy(), Eelec=Eelec.detach().numpy(), Enuc=Enuc.detach().numpy(),
             Eiso=Eiso.detach().numpy(), EnucAB=EnucAB.detach().numpy(), notconverged=notconv.numpy(),
             trace=np.array(trace, dtype=np.float64), ref_seconds=dt)
    if case['keep_w']:
        g['w'] = w.numpy()
    else:
        g['w_sum'] = w.sum(dim=(1, 2)).numpy()
        g['w_00'] = w[:, 0, 0].numpy()
    return g


def compare_oracle(case, g):
    sys.path.insert(0, os.path.join(ROOT, 'oracle'))
    import seqm_oracle as orc
    trace = []
    o = orc.run(torch.as_tensor(case['species']), torch.as_tensor(case['coords']), method=case['method'],
                scf_eps=case['eps'], converger=tuple(case['conv']), sp2=tuple(case['sp2']), trace=trace)
    d = {}
    for k in ('mask', 'maskd', 'idxi', 'idxj', 'ni', 'nj', 'pair_molid', 'atom_molid', 'Z'):
        d[k] = int((o['ps'][k].numpy() != g[k]).sum())
    for k in ('xij', 'rij'):
        d[k] = float(np.abs(o['ps'][k].detach().numpy() - g[k]).max())
    for k in ('M', 'F', 'P', 'force', 'Etot', 'Eelec', 'Hf', 'EnucAB'):
        d[k] = float(np.abs(o[k].numpy() - g[k]).max())
    if 'w' in g:
        d['w'] = float(np.abs(o['w'].numpy() - g['w']).max())
    d['trace_ref'] = [int(t[2]) for t in g['trace']]
    d['trace_orc'] = [t[2] for t in trace]
    return d


def main():
    os.makedirs(OUT, exist_ok=True)
    seqm = load_reference()
    only = sys.argv[1:]
    for case in cases():
        if only and not any(s in case['name'] for s in only):
            continue
        g = run_reference(seqm, case)
        np.savez_compressed(os.path.join(OUT, case['name'] + '.npz'), **g)
        d = compare_oracle(case, g)
        same = d['trace_ref'] == d['trace_orc']
        print(case['name'], 'ref %.2fs' % g['ref_seconds'], 'iters', len(d['trace_ref']), 'trace_same', same,
              {k: v for k, v in d.items() if not k.startswith('trace')})
        if not same:
            print('   ref', d['trace_ref'], '\n   orc', d['trace_orc'])


if __name__ == '__main__':
    main()
