"""Is a Force call reproducible bit for bit (same process, same inputs), with and without the sharded SP2 form?"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
torch.set_default_dtype(torch.float64)
from pyseqm_b200 import _ops, synth
from pyseqm_b200.basics import Force
from pyseqm_b200.seqm_functions.constants import Constants
dev = torch.device('cuda:0')
sp, xyz = synth.c60(3, seed=2)
par = dict(method='PM3', elements=[0, 6], scf_eps=1.0e-6, scf_converger=[2, 0.0], sp2=[True, 1.0e-5], learned=[],
           pair_outer_cutoff=1.0e10)
species, coords = torch.as_tensor(sp).to(dev), torch.as_tensor(xyz).to(dev)
const = Constants().to(dev)
def run(ranks, fresh):
    _ops.sp2_sharding(None, ranks)
    try:
        fm = Force(par).to(dev) if fresh else run.fm
        run.fm = fm
        o = fm(const, coords.clone(), species)
        st = fm.energy.last_state
        return o[2].detach().clone(), o[0].clone(), st['F'].clone(), st['niter'].clone()
    finally:
        _ops.sp2_sharding(None, 0)
a = run(0, True); b = run(0, False); c = run(0, True); d = run(3, True); e = run(3, False)
for nm, x in (('same module again', b), ('fresh module', c), ('sharded(3) fresh', d), ('sharded(3) again', e)):
    print('%-20s dE %.3e dForce %.3e dF %.3e niter %s' % (nm, float((a[0] - x[0]).abs().max()), float((a[1] - x[1]).abs().max()),
                                                           float((a[2] - x[2]).abs().max()), x[3].tolist()))
