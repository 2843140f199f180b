"""Config-4 molecules with more than 32 packed orbitals (the CTA Jacobi path): sweep time under the eigensolver
variants (SEQM_EIGH_CFG bits, SEQM_EIGH_TOL) and the agreement of the energies / forces with the default build."""
import ctypes, os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
torch.set_default_dtype(torch.float64)
from pyseqm_b200 import _lib, synth
from pyseqm_b200.basics import Force
from pyseqm_b200.batching import BucketedBatch, packed_sizes
from pyseqm_b200.seqm_functions.constants import Constants
dev = torch.device('cuda:0')
par = dict(method='MNDO', scf_eps=1.0e-6, scf_converger=[2, 0.0], sp2=[False], elements=[0, 1, 6, 7, 8], learned=[],
           pair_outer_cutoff=1.0e10)
sp, xyz = synth.mixed_organics_sweep(int(os.environ.get('NMOL', '16384')), seed=0)
species, coords = torch.as_tensor(sp).to(dev), torch.as_tensor(xyz).to(dev)
big = packed_sizes(species) > 32
species, coords = species[big].contiguous(), coords[big].contiguous()
const = Constants().to(dev)
L = _lib.lib()
ref = None
for cfg, tol in ((0, 0.0), (1, 0.0), (2, 0.0), (3, 0.0), (0, 1e-18), (3, 1e-18), (3, 1e-16)):
    L.seqm_eigh_block_config(cfg)
    L.seqm_eigh_tolerance(tol)
    fm = Force(par).to(dev)
    bb = BucketedBatch(species)
    for _ in range(2):
        f, E, Hf, nc = bb.run(fm, const, coords)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        f, E, Hf, nc = bb.run(fm, const, coords)
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) / 3 * 1e3
    if ref is None:
        ref = (E.clone(), f.clone())
    print('cfg %d tol %-7g  %8.2f ms per sweep of %d molecules  notconverged %d  max|dE| %.2e  max|dF| %.2e' %
          (cfg, tol, ms, species.shape[0], int(nc.sum()), float((E - ref[0]).abs().max()), float((f - ref[1]).abs().max())))
