"""MD drivers on the GPU path against golden trajectories of the unmodified reference
(tools/make_golden_md.py): BOMD (velocity Verlet around Force) and XL-BOMD (EnergyXL/ForceXL, SP2 and eigh)."""
import contextlib
import io
import os

import numpy as np
import pytest
import torch

from conftest import ROOT

pytestmark = pytest.mark.gpu
torch.set_default_dtype(torch.float64)
GMD = os.path.join(ROOT, 'tests', 'golden_md')


@pytest.mark.parametrize('tag,mode', [('bomd_eigh', 'autograd'), ('xlbomd_sp2', 'fused'), ('xlbomd_sp2', 'graph'),
                                      ('xlbomd_sp2', 'autograd'), ('xlbomd_eigh', 'fused'), ('xlbomd_eigh', 'graph'),
                                      ('xlbomd_eigh', 'autograd'), ('xlbomd_c60_sp2', 'fused'), ('xlbomd_c60_sp2', 'graph'),
                                      ('xlbomd_c60_sp2', 'autograd')])
def test_md_trajectory(tag, mode):
    """mode (XL-BOMD only): 'fused' = the step as a fixed sequence of library launches (_xlstep.XLStep, the default),
    'graph' = the same captured in one CUDA graph, 'autograd' = the reference-shaped form (EnergyXL + torch.autograd)"""
    from pyseqm_b200.MolecularDynamics import Molecular_Dynamics_Basic
    from pyseqm_b200.XLBOMD import XL_BOMD
    from pyseqm_b200.seqm_functions.constants import Constants
    dev = torch.device('cuda:0')
    g = np.load(os.path.join(GMD, tag + '.npz'))
    sp2 = g['sp2'].tolist()
    par = dict(method=str(g['method']) if 'method' in g else 'AM1', scf_eps=float(g['eps']) if 'eps' in g else 1e-8,
               scf_converger=[2, 0.0], sp2=[bool(sp2[0])] + sp2[1:],
               elements=[0, 1, 6, 8], learned=[], pair_outer_cutoff=1e10, xl_fused=(mode != 'autograd'),
               xl_cuda_graph=(mode == 'graph'))
    out = {'molid': [0, 1], 'thermo': 1, 'dump': 10 ** 9, 'prefix': '/tmp/seqm_b200_md'}
    const = Constants().to(dev)
    if tag.startswith('xl'):
        md = XL_BOMD(par, timestep=1.0, k=int(g['k']), output=out).to(dev)
    else:
        md = Molecular_Dynamics_Basic(par, timestep=1.0, output=out).to(dev)
    c = torch.as_tensor(g['coords0']).to(dev)
    v = torch.as_tensor(g['vel0']).to(dev)
    species = torch.as_tensor(g['species']).to(dev)
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        res = md.run(const, int(g['steps']), c, v, species)
    thermo = []
    for line in buf.getvalue().splitlines():
        t = line.split()
        if len(t) == 15 and t[0].isdigit():
            thermo.append([float(x) for x in t[1:]])
    thermo = np.array(thermo)
    assert thermo.shape == g['thermo'].shape
    # printed with 7 significant digits
    assert np.abs(thermo - g['thermo']).max() < 2e-5 * max(1.0, np.abs(g['thermo']).max())
    assert np.abs(res[0].detach().cpu().numpy() - g['coords']).max() < 1e-7
    assert np.abs(res[1].detach().cpu().numpy() - g['vel']).max() < 1e-7
    if tag.startswith('xl'):
        xs = getattr(md, '_xlstep', None)
        assert (xs is not None) == (mode != 'autograd')
        if mode == 'graph':
            assert xs.graph is not None and xs.nsteps == int(g['steps'])


def test_standalone_ops_match_oracle():
    """the reference-signature functions hcore / fock / SP2 / sym_eig_trunc / pack / unpack"""
    import seqm_oracle as orc
    from pyseqm_b200 import synth
    from pyseqm_b200.basics import Pack_Parameters, Parser
    from pyseqm_b200.seqm_functions.constants import Constants
    from pyseqm_b200.seqm_functions.diag import sym_eig_trunc
    from pyseqm_b200.seqm_functions.fock import fock
    from pyseqm_b200.seqm_functions.hcore import hcore
    from pyseqm_b200.seqm_functions.pack import pack, unpack
    from pyseqm_b200.seqm_functions.SP2 import SP2
    dev = torch.device('cuda:0')
    sp_, xyz = synth.mixed_organics(6, seed=11, min_atoms=5, max_atoms=12)
    par = dict(method='PM3', scf_eps=1e-6, scf_converger=[2, 0.0], sp2=[False], elements=[0, 1, 6, 7, 8], learned=[],
               pair_outer_cutoff=1e10)
    const = Constants().to(dev)
    species, coords = torch.as_tensor(sp_).to(dev), torch.as_tensor(xyz).to(dev)
    nmol, molsize, nHeavy, nHydro, nocc, Z, maskd, atom_molid, mask, pair_molid, ni, nj, idxi, idxj, xij, rij = \
        Parser(par)(const, species, coords)
    p = Pack_Parameters(par).to(dev)(Z, learned_params=dict())
    beta = torch.stack((p['beta_s'], p['beta_p']), dim=1)
    M, w = hcore(const, nmol, molsize, maskd, mask, idxi, idxj, ni, nj, xij, rij, Z, p['zeta_s'], p['zeta_p'],
                 p['U_ss'], p['U_pp'], p['g_ss'], p['g_pp'], p['g_p2'], p['h_sp'], beta)
    ref = orc.run(torch.as_tensor(sp_), torch.as_tensor(xyz), method='PM3', scf_eps=1e-6, forces=False)
    assert float((M.cpu() - ref['M']).abs().max()) < 1e-9
    assert float((w.cpu() - ref['w']).abs().max()) < 1e-9
    Pd = ref['P'].to(dev)
    F = fock(nmol, molsize, Pd, M, maskd, mask, idxi, idxj, w, p['g_ss'], p['g_pp'], p['g_sp'], p['g_p2'], p['h_sp'])
    assert float((F.cpu() - ref['F']).abs().max()) < 1e-8
    e, Pn, v = sym_eig_trunc(F, nHeavy, nHydro, nocc)
    assert float((Pn.cpu() - ref['P']).abs().max()) < 1e-4         # ref P is converged to sqrt(eps) only
    Fp = pack(F, nHeavy, nHydro)
    assert float((Fp.cpu() - orc.pack(ref['F'], ref['ps']['nHeavy'], ref['ps']['nHydro'])).abs().max()) < 1e-8
    assert float((unpack(Fp, nHeavy, nHydro, 4 * molsize) - F).abs().max()) < 1e-12
    # SP2 on one molecule's packed matrix
    m0 = int(4 * nHeavy[0] + nHydro[0])
    D = SP2(Fp[:1, :m0, :m0].contiguous(), nocc[:1], 1e-7)
    Dref, _ = orc.sp2_density(Fp[:1, :m0, :m0].cpu(), nocc[:1].cpu(), 1e-7)
    assert float((D.cpu() - Dref).abs().max()) < 1e-9
