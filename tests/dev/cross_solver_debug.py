"""per-molecule differences between the two eigensolver variants through a whole Force call (debug helper of
tests/test_gpu_parity.py::test_ring_jacobi_against_two_sided_jacobi)"""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
torch.set_default_dtype(torch.float64)
from pyseqm_b200 import _lib
from pyseqm_b200.basics import Force
from pyseqm_b200.seqm_functions.constants import Constants
import test_gpu_parity as T
dev = torch.device('cuda:0')
name = sys.argv[1] if len(sys.argv) > 1 else 'small_PM3_eigh_c2'
g = T._load(name)
const = Constants().to(dev)
res = {}
for variant in (1, 0):
    _lib.lib().seqm_eigh_warp_variant(variant)
    fo = Force(T._params(g)).to(dev)
    out = fo(const, T._t(g["coords"], dev), T._t(g["species"], dev))
    torch.cuda.synchronize()
    res[variant] = (out[0].clone().cpu(), out[3].detach().clone().cpu(), fo.energy.last_state['niter'].clone().cpu())
_lib.lib().seqm_eigh_warp_variant(1)
fref = torch.as_tensor(g['force']) if 'force' in g.files else None
print('species', g['species'].tolist())
for b in range(res[1][0].shape[0]):
    line = 'mol %d  dF %.3e  dHf %.3e  niter %d / %d' % (b, float((res[1][0][b] - res[0][0][b]).abs().max()),
                                                       float((res[1][1][b] - res[0][1][b]).abs()),
                                                       int(res[1][2][b]), int(res[0][2][b]))
    if fref is not None:
        line += '  vs ref: ring %.3e two-sided %.3e' % (float((res[1][0][b] - fref[b]).abs().max()),
                                                        float((res[0][0][b] - fref[b]).abs().max()))
    print(line)
