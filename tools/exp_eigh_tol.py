"""Effect of the Jacobi stopping tolerance on the converged SCF (1024 benzenes, bench workload): time per Force
call and the deviation of P / Etot / forces from the tightest setting."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
torch.set_default_dtype(torch.float64)
from pyseqm_b200 import _lib, synth  # noqa: E402
from pyseqm_b200.basics import Force  # noqa: E402
from pyseqm_b200.seqm_functions.constants import Constants  # noqa: E402

dev = torch.device('cuda:0')
par = dict(method='AM1', scf_eps=1.0e-6, scf_converger=[2, 0.0], sp2=[False], elements=[0, 1, 6], learned=[],
           pair_outer_cutoff=1.0e10)
sp, xyz = synth.benzene(1024, seed=0)
species = torch.as_tensor(sp).to(dev)
coords = torch.as_tensor(xyz).to(dev)
const = Constants().to(dev)
f = Force(par).to(dev)
ref = None
for tol in (1e-30, 1e-27, 1e-24, 1e-21, 1e-18, 1e-16, 1e-14, 1e-12):
    _lib.lib().seqm_eigh_tolerance(tol)
    for _ in range(3):
        out = f(const, coords.clone(), species)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5):
        out = f(const, coords.clone(), species)
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) / 5 * 1e3
    st = f.energy.last_state
    cur = (out[1].clone(), out[2].detach().clone(), out[0].clone())
    if ref is None:
        ref = cur
    print('tol %.0e  %.3f ms  iters %d  dP %.2e  dE %.2e  dForce %.2e  niter-sum %d' % (
        tol, ms, st['iterations'], float((cur[0] - ref[0]).abs().max()), float((cur[1] - ref[1]).abs().max()),
        float((cur[2] - ref[2]).abs().max()), int(st['niter'].sum())))
