"""GraphedForce (one Force call replayed as a CUDA graph) against the eager Force and the reference fixture."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN

pytestmark = pytest.mark.gpu
torch.set_default_dtype(torch.float64)


def test_graphed_force_matches_eager_and_reference():
    from pyseqm_b200.basics import Force, GraphedForce
    from pyseqm_b200.seqm_functions.constants import Constants
    dev = torch.device('cuda:0')
    g = np.load(os.path.join(GOLDEN, 'benzene8_AM1_eigh_c2.npz'))
    par = dict(method='AM1', scf_eps=float(g['eps']), scf_converger=[2, 0.0], sp2=[False], elements=[0, 1, 6], learned=[],
               pair_outer_cutoff=1.0e10)
    const = Constants().to(dev)
    species = torch.as_tensor(g['species']).to(dev)
    coords = torch.as_tensor(g['coords']).to(dev)
    gf = GraphedForce(par).to(dev)
    out = gf(const, coords, species)
    torch.cuda.synchronize()
    assert float((out[2].cpu() - torch.as_tensor(g['Etot'])).abs().max()) < 1e-6
    assert float((out[0].cpu() - torch.as_tensor(g['force'])).abs().max()) < 1e-5
    assert not bool(out[10].any())
    eager = Force(par).to(dev)
    rng = torch.Generator().manual_seed(3)
    for _ in range(3):      # replays on new coordinates: identical to the eager call, outputs are the same buffers
        x = coords + 0.02 * torch.randn(coords.shape, generator=rng).to(dev) * (species > 0).unsqueeze(-1)
        o1 = gf(const, x, species)
        assert o1[0].data_ptr() == out[0].data_ptr()
        o2 = eager(const, x.clone(), species)
        torch.cuda.synchronize()
        assert float((o1[2] - o2[2]).abs().max()) < 1e-10
        assert float((o1[0] - o2[0]).abs().max()) < 1e-10
        assert float((o1[1] - o2[1]).abs().max()) < 1e-10
    # another species tensor (half the batch): recapture
    sp2, x2 = species[:4].clone(), coords[:4].clone()
    o3 = gf(const, x2, sp2)
    o4 = eager(const, x2.clone(), sp2)
    torch.cuda.synchronize()
    assert o3[0].shape == o4[0].shape
    assert float((o3[2] - o4[2]).abs().max()) < 1e-10
    assert float((o3[0] - o4[0]).abs().max()) < 1e-10
