"""Short driver for ncu captures at config-3 size: a few fused XL-BOMD steps of 256 C60 (sp2_step_kernel, pair kernels,
xl_extrapolate_kernel, verlet / energy kernels) and one eigh density solve of 16 C60 (eigh_cluster_kernel)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
torch.set_default_dtype(torch.float64)
from pyseqm_b200 import _ops, synth
from pyseqm_b200.XLBOMD import XL_BOMD
from pyseqm_b200.seqm_functions.constants import Constants
dev = torch.device('cuda:0')
const = Constants().to(dev)
nmol = int(os.environ.get('NMOL', '256'))
par = dict(method='PM3', scf_eps=1.0e-6, scf_converger=[2, 0.0], sp2=[True, 1.0e-5], elements=[0, 6], learned=[],
           pair_outer_cutoff=1.0e10)
out = {'molid': [0], 'thermo': 10 ** 9, 'dump': 10 ** 9, 'prefix': '/tmp/seqm_b200_ncu'}
sp, xyz = synth.c60(nmol, seed=0)
species, coords = torch.as_tensor(sp).to(dev), torch.as_tensor(xyz).to(dev)
md = XL_BOMD(par, timestep=1.0, k=5, output=out).to(dev)
torch.manual_seed(0)
vel = md.initialize_velocity(const, coords, species, Temp=300.0)
mass = md._masses(const, species)
acc, D = md.initialize(const, mass, coords, species)
Pt = D.unsqueeze(0).expand((md.m,) + D.shape).clone()
P = D.clone()
state = [coords, vel, acc, D, P, Pt]
for i in range(3):
    c, v, a, d, p_, pt, Hf, force = md.one_step(const, i, mass, *state[:2], species, *state[2:])
    state = [c, v, a, d, p_, pt]
torch.cuda.synchronize()
F = md._xlstep.F[:16].contiguous()
topo = md._xlstep.topo
Pe = _ops.eigh_density(F, topo.nheavy[:16].contiguous(), topo.nhydro[:16].contiguous(), topo.nocc[:16].contiguous(), 60, 240)
torch.cuda.synchronize()
print('ok', float(Hf.sum()), float(Pe.diagonal(dim1=1, dim2=2).sum()))
