"""Cycle accounting of the one-launch SCF kernel (instrumented build, molecule 0 of the 1024-benzene workload)."""
import ctypes, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
torch.set_default_dtype(torch.float64)
from pyseqm_b200 import _lib, synth
from pyseqm_b200.basics import Force
from pyseqm_b200.seqm_functions.constants import Constants
dev = torch.device('cuda:0')
par = dict(method='AM1', scf_eps=1.0e-6, scf_converger=[2, 0.0], sp2=[False], elements=[0, 1, 6], learned=[],
           pair_outer_cutoff=1.0e10)
sp, xyz = synth.benzene(int(os.environ.get('NMOL', '1024')), seed=0)
species, coords = torch.as_tensor(sp).to(dev), torch.as_tensor(xyz).to(dev)
const = Constants().to(dev)
f = Force(par).to(dev)
L = _lib.lib()
L.seqm_debug_phase_ticks.argtypes = [ctypes.POINTER(ctypes.c_longlong), ctypes.c_int]
buf = (ctypes.c_longlong * 16)()
for _ in range(2):
    f(const, coords.clone(), species)
L.seqm_debug_phase_ticks(buf, 1)
f(const, coords.clone(), species)
L.seqm_debug_phase_ticks(buf, 1)
names = ['setup (schedule, P0)', 'fock: core Hamiltonian init', 'fock: pairs', 'fock: atoms (diagonal blocks)', 'fock: energy',
         'pulay store', 'pulay extrapolate (+loop control)', 'eigh: load + warm-start transform', 'eigh: sweeps',
         'eigh: density P = 2 W W^T', 'mix', 'loop control before fock', 'eigh: eigenvalues + ranking', 'eigh: store V / W']
tot = sum(buf[i] for i in range(16))
st = f.energy.last_state
print('molecule 0: %d iterations, total %.0f kcycles' % (int(st['niter'][0]), tot / 1e3))
for i, n in enumerate(names):
    print('  %-36s %9.1f kcycles  %5.1f %%' % (n, buf[i] / 1e3, 100.0 * buf[i] / tot))
