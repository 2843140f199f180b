import sys, torch
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
torch.set_default_dtype(torch.float64)
from pyseqm_b200 import synth
from pyseqm_b200.XLBOMD import XL_BOMD
from pyseqm_b200.seqm_functions.constants import Constants
dev = torch.device('cuda:0')
par = dict(method='PM3', scf_eps=1.0e-6, scf_converger=[2, 0.0], sp2=[True, 1.0e-5], elements=[0, 6], learned=[], pair_outer_cutoff=1.0e10)
sp, xyz = synth.c60(4, seed=0)
species, coords = torch.as_tensor(sp).to(dev), torch.as_tensor(xyz).to(dev)
const = Constants().to(dev)
out = {'molid': [0], 'thermo': 10 ** 9, 'dump': 10 ** 9, 'prefix': '/tmp/x'}
md = XL_BOMD(par, timestep=1.0, k=5, output=out).to(dev)
torch.manual_seed(0)
c = coords.clone(); vel = md.initialize_velocity(const, c, species, Temp=300.0)
mass = md._masses(const, species)
acc, D = md.initialize(const, mass, c, species)
Pt = D.unsqueeze(0).expand((md.m,) + D.shape).clone(); Pd = D.clone()
for i in range(30):
    c, vel, acc, D, Pd, Pt, Hf, frc = md.one_step(const, i, mass, c, vel, species, acc, D, Pd, Pt)
    ek, T = md.kinetic_energy(const, mass, species, vel)
    print(i, 'Ek %.4f  Hf %.4f  tot %.5f  T %.1f' % (float(ek[0]), float(Hf[0]), float(ek[0] + Hf[0]), float(T[0])))
