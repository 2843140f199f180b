"""The oracle (oracle/seqm_oracle.py) against the golden vectors the unmodified reference produced
(tools/make_golden.py).  This is what pins the oracle; it needs neither the reference nor a GPU."""
import os

import numpy as np
import pytest
import torch

import seqm_oracle as orc
from conftest import GOLDEN, golden_names

# molecules whose DIIS error vectors vanish by symmetry (N2): the Pulay coefficients are then ratios of
# rounding noise and even two builds of the reference disagree at 1e-6 in P; energies still agree.
NOISE_DIIS = {'small_PM3_eigh_c2': [1], 'small_MNDO_eigh_c2': [1]}


@pytest.mark.parametrize('name', golden_names())
def test_oracle_matches_reference(name):
    g = np.load(os.path.join(GOLDEN, name + '.npz'))
    trace = []
    sp2 = tuple(g['sp2'].tolist())
    sp2 = (bool(sp2[0]),) + tuple(sp2[1:])
    o = orc.run(torch.as_tensor(g['species']), torch.as_tensor(g['coords']), method=str(g['method']),
                scf_eps=float(g['eps']), converger=tuple(g['conv'].tolist()), sp2=sp2, trace=trace)
    # Parser tensors: bit exact
    for k in ('mask', 'maskd', 'idxi', 'idxj', 'ni', 'nj', 'pair_molid', 'atom_molid', 'Z', 'nHeavy', 'nHydro',
              'nocc'):
        assert np.array_equal(o['ps'][k].numpy(), g[k]), k
    for k in ('xij', 'rij'):
        assert np.array_equal(o['ps'][k].detach().numpy(), g[k]), k
    # direct formulas
    assert np.abs(o['M'].numpy() - g['M']).max() < 1e-11
    assert np.abs(o['EnucAB'].numpy() - g['EnucAB']).max() < 1e-11
    if 'w' in g:
        assert np.abs(o['w'].numpy() - g['w']).max() < 1e-12
    else:
        assert np.abs(o['w'].sum(dim=(1, 2)).numpy() - g['w_sum']).max() < 1e-11
    # iteration sequence (number of unconverged molecules after each iteration)
    assert [t[2] for t in trace] == [int(t[2]) for t in g['trace']]
    keep = np.ones(g['species'].shape[0], dtype=bool)
    keep[NOISE_DIIS.get(name, [])] = False
    assert np.abs(o['Etot'].numpy() - g['Etot']).max() < 1e-9
    assert np.abs(o['Eelec'].numpy() - g['Eelec']).max() < 1e-9
    assert np.abs(o['Hf'].numpy() - g['Hf']).max() < 1e-9
    assert np.abs(o['P'].numpy() - g['P'])[keep].max() < 1e-8
    assert np.abs(o['F'].numpy() - g['F'])[keep].max() < 1e-8
    assert np.abs(o['force'].numpy() - g['force'])[keep].max() < 1e-7
    assert np.array_equal(o['notconverged'].numpy(), g['notconverged'])
