"""Which of (Jacobi stopping tolerance, device pair geometry) moves the converged density of a golden case."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pyseqm_b200 import _lib  # noqa: E402
from pyseqm_b200.basics import Force  # noqa: E402
from pyseqm_b200.seqm_functions.constants import Constants  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else 'mixed12_MNDO_eigh_c2'
g = dict(np.load(os.path.join(os.path.dirname(__file__), '..', 'tests', 'golden', name + '.npz'), allow_pickle=True))
torch.set_default_dtype(torch.float64)
dev = torch.device('cuda:0')
const = Constants().to(dev)
species = torch.as_tensor(g['species']).to(dev)
conv = g['conv'].tolist()
conv[0] = int(conv[0])
sp2 = g['sp2'].tolist()
sp2 = [bool(sp2[0])] + sp2[1:]
for tol in (1e-18, 1e-22, 1e-26, 1e-30):
    for cutoff in (1e10, 1e4):
        _lib.lib().seqm_eigh_tolerance(tol)
        sp = dict(method=str(g['method']), scf_eps=float(g['eps']), scf_converger=conv, sp2=sp2,
                  elements=[0, 1, 6, 7, 8], learned=[], pair_outer_cutoff=cutoff)
        out = Force(sp).to(dev)(const, torch.as_tensor(g['coords']).to(dev).clone(), species)
        P = out[1].cpu()
        dP = (P - torch.as_tensor(g['P'])).abs().amax(dim=(1, 2))
        dF = (out[0].cpu() - torch.as_tensor(g['force'])).abs().amax(dim=(1, 2))
        print('tol %.0e cutoff %.0e  max dP %.3e (mol %d)  max dForce %.3e' % (tol, cutoff, float(dP.max()),
                                                                             int(dP.argmax()), float(dF.max())))
