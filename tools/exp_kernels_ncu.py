"""Short driver for ncu captures of kernels that bench.py reaches only inside long runs: one Force call on 128 C60
(pair_grad_xx_kernel, sp2_step_kernel) and one on 2048 organics with more than 32 packed orbitals (eigh_density_kernel)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
torch.set_default_dtype(torch.float64)
from pyseqm_b200 import synth
from pyseqm_b200.basics import Force
from pyseqm_b200.batching import packed_sizes
from pyseqm_b200.seqm_functions.constants import Constants
dev = torch.device('cuda:0')
const = Constants().to(dev)
sp, xyz = synth.c60(128, seed=0)
par = dict(method='PM3', scf_eps=1.0e-6, scf_converger=[2, 0.0], sp2=[True, 1.0e-5], elements=[0, 6], learned=[],
           pair_outer_cutoff=1.0e10)
out = Force(par).to(dev)(const, torch.as_tensor(xyz).to(dev), torch.as_tensor(sp).to(dev))
sp, xyz = synth.mixed_organics_sweep(8192, seed=0)
species, coords = torch.as_tensor(sp).to(dev), torch.as_tensor(xyz).to(dev)
m = packed_sizes(species)
big = (m > 40) & (m <= 48)
species, coords = species[big][:2048].contiguous(), coords[big][:2048].contiguous()
par = dict(method='MNDO', scf_eps=1.0e-6, scf_converger=[2, 0.0], sp2=[False], elements=[0, 1, 6, 7, 8], learned=[],
           pair_outer_cutoff=1.0e10)
out2 = Force(par).to(dev)(const, coords, species)
torch.cuda.synchronize()
print('ok', float(out[2].sum()), float(out2[2].sum()), int(species.shape[0]))
