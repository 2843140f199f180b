import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))      # development aid (uses the oracle as checker): lives under tests/
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'oracle'))
torch.set_default_dtype(torch.float64)
from pyseqm_b200 import _engine, _ops
from pyseqm_b200.basics import Pack_Parameters, Parser
from pyseqm_b200.seqm_functions.constants import Constants
import seqm_oracle as orc
dev = torch.device('cuda:0')
g = np.load(os.path.join(ROOT, 'tests/golden/c60_PM3_sp2_c2.npz'))
sp = dict(method='PM3', scf_eps=1e-6, scf_converger=[2, 0.0], sp2=[True, 1e-5], elements=[0, 1, 6, 7, 8], learned=[], pair_outer_cutoff=1e10)
const = Constants().to(dev)
species, coords = torch.as_tensor(g['species']).to(dev), torch.as_tensor(g['coords']).to(dev)
out = Parser(sp)(const, species, coords)
nmol, molsize, nHeavy, nHydro, nocc, Z, maskd, atom_molid, mask, pair_molid, ni, nj, idxi, idxj, xij, rij = out
par = Pack_Parameters(sp).to(dev)(Z, learned_params=dict())
st = _engine.build(const, nmol, molsize, nHeavy, nHydro, nocc, Z, maskd, idxi, idxj, xij, rij, par, 'PM3', run_scf=False)
ap = st['atompar'].cpu().numpy().view(np.float64).reshape(-1, 28)
tab = orc.load_params('PM3'); Zc = Z.cpu()
p = {k: v[Zc] for k, v in tab.items()}
dd, qq, r0, r1, r2 = orc.atom_multipoles(Zc, p['zeta_s'], p['zeta_p'], p['g_ss'], p['g_pp'], p['g_p2'], p['h_sp'])
print('multipole rel diffs', [float(np.abs(ap[:, 4 + i] / x.numpy() - 1).max()) for i, x in enumerate((dd, qq, r0, r1, r2))])
dM = (st['M'].cpu() - torch.as_tensor(g['M'])).abs()
blk = dM.reshape(dM.shape[0], -1).max(1)[0]
b = int(blk.argmax()); S = molsize
print('worst block', b, 'slots', (b % (S * S)) // S, b % S, 'diff', dM[b])
print('M block', st['M'][b].cpu(), '\nref', torch.as_tensor(g['M'])[b])
diag = torch.as_tensor(g['maskd'])
print('max diag-block diff', float(dM[diag].max()), 'max offdiag diff', float(dM[torch.as_tensor(g['mask'])].max()))
w = _ops.w_expand(st['topo'], st['w']).cpu()
# oracle w with reference geometry
ps = orc.parse(torch.as_tensor(g['species']), torch.as_tensor(g['coords']))
wo, e1b, e2a = orc.two_center_integrals(ps['ni'], ps['nj'], ps['xij'], ps['rij'], dd, qq, r0, r1, r2, ps['idxi'], ps['idxj'])
print('w max diff vs oracle', float((w - wo).abs().max()), 'rel', float(((w - wo).abs() / (wo.abs() + 1e-3)).max()))
pm = int(dM[torch.as_tensor(g['mask'])].reshape(-1, 16).max(1)[0].argmax())
print('worst pair', pm, 'rij', float(ps['rij'][pm]), 'xij', ps['xij'][pm])
