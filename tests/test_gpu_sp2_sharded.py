"""Tile-sharded SP2 (BASELINE.json configs[4]: one large system, the X @ X of reference SP2.py:47 split over GPUs).
 * emulated ranks on one GPU: the tile ranges of 1, 2, 3 and 8 ranks go through the record / unpack path one after the
   other; result and iteration count must be BIT-IDENTICAL to seqm_sp2_density (same tile arithmetic, same trace sums)
 * the same through Force (seqm_scf hook)
 * real NCCL, world size 2: tests/multi_gpu/run_sp2_nccl.py under torchrun (skipped with fewer than 2 GPUs)
"""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
torch.set_default_dtype(torch.float64)
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _system(kind, dev):
    from pyseqm_b200 import synth
    if kind == 'c60x3':
        sp, xyz = synth.c60(3, seed=2)
        par = dict(method='PM3', elements=[0, 6])
    else:
        sp, xyz = synth.water_cluster(60, seed=3)
        par = dict(method='PM3', elements=[0, 1, 8])
    par.update(scf_eps=1.0e-6, scf_converger=[2, 0.0], sp2=[True, 1.0e-5], learned=[], pair_outer_cutoff=1.0e10)
    return par, torch.as_tensor(sp).to(dev), torch.as_tensor(xyz).to(dev)


@pytest.mark.parametrize('kind', ['c60x3', 'water60'])
def test_emulated_ranks_bit_identical(kind):
    from pyseqm_b200 import _ops
    from pyseqm_b200.basics import Force
    from pyseqm_b200.seqm_functions.constants import Constants
    dev = torch.device('cuda:0')
    par, species, coords = _system(kind, dev)
    const = Constants().to(dev)
    fm = Force(par).to(dev)
    out = fm(const, coords.clone(), species)
    st = fm.energy.last_state
    topo, F = st['topo'], st['F']
    molsize = species.shape[1]
    D0, n0 = _ops.sp2_density(F, topo.nheavy, topo.nhydro, topo.nocc, molsize, topo.mmax, 1e-5, want_niter=True)
    for ranks in (1, 2, 3, 8):
        for nsub in (0, 1, 4):
            _ops.lib().seqm_sp2_nccl_subchunks(nsub)
            D1, n1 = _ops.sp2_density_sharded(F, topo.nheavy, topo.nhydro, topo.nocc, molsize, topo.mmax, 1e-5,
                                              comm=None, ranks=ranks, want_niter=True)
            assert torch.equal(n0, n1), (ranks, nsub)
            assert torch.equal(D0, D1), (ranks, nsub, float((D0 - D1).abs().max()))
    _ops.lib().seqm_sp2_nccl_subchunks(0)
    # through the SCF driver
    try:
        _ops.sp2_sharding(None, 3)
        out2 = Force(par).to(dev)(const, coords.clone(), species)
    finally:
        _ops.sp2_sharding(None, 0)
    # (the SCF's convergence test works on energy differences of ~1e-7 eV, so two runs agree to the SCF threshold)
    assert float((out[2] - out2[2]).abs().max()) < 1e-6 and float((out[0] - out2[0]).abs().max()) < 1e-5


def test_nccl_world2():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2', '--master-addr',
           '127.0.0.1', '--master-port', '29631', os.path.join(ROOT, 'tests', 'multi_gpu', 'run_sp2_nccl.py')]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert 'SP2-NCCL-OK' in r.stdout
