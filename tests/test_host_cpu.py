"""CPU-side checks of the host layer: the Parser is bit-exact against the reference's golden tensors, the
C-ABI library loads and exports every symbol include/seqm_b200.h declares, and the product refuses to
compute without CUDA (no fallback)."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from conftest import GOLDEN, ROOT, golden_names

torch.set_default_dtype(torch.float64)


@pytest.mark.parametrize('name', golden_names())
def test_parser_bit_exact(name):
    from pyseqm_b200.basics import Parser
    from pyseqm_b200.seqm_functions.constants import Constants
    g = np.load(os.path.join(GOLDEN, name + '.npz'))
    sp = dict(pair_outer_cutoff=1.0e10, elements=[0, 1, 6, 7, 8])
    out = Parser(sp)(Constants(), torch.as_tensor(g['species']), torch.as_tensor(g['coords']))
    names = ['nmol', 'molsize', 'nHeavy', 'nHydro', 'nocc', 'Z', 'maskd', 'atom_molid', 'mask', 'pair_molid', 'ni',
             'nj', 'idxi', 'idxj', 'xij', 'rij']
    for k, v in zip(names, out):
        if k in ('nmol', 'molsize'):
            continue
        assert v.dtype == (torch.float64 if k in ('xij', 'rij') else torch.int64), k
        assert np.array_equal(v.numpy(), g[k]), k


def test_parser_cutoff_and_charges():
    from pyseqm_b200.basics import Parser
    from pyseqm_b200.seqm_functions.constants import Constants
    import seqm_oracle as orc
    from pyseqm_b200 import synth
    sp_, xyz = synth.water_cluster(5, seed=1)
    species, coords = torch.as_tensor(sp_), torch.as_tensor(xyz)
    out = Parser(dict(pair_outer_cutoff=4.0, elements=[0, 1, 8]))(Constants(), species, coords)
    ref = orc.parse(species, coords, cutoff=4.0)
    assert np.array_equal(out[12].numpy(), ref['idxi'].numpy()) and np.array_equal(out[13].numpy(), ref['idxj'].numpy())
    assert np.array_equal(out[15].numpy(), ref['rij'].numpy())
    with pytest.raises(ValueError):
        Parser(dict(pair_outer_cutoff=1e10, elements=[0, 1, 8]))(Constants(), species, coords,
                                                                 charges=torch.tensor([1]))


def test_library_exports_header_symbols():
    from pyseqm_b200 import _lib
    hdr = open(os.path.join(ROOT, 'include', 'seqm_b200.h')).read()
    declared = set(re.findall(r'\b(seqm_[a-z0-9_]+)\s*\(', hdr))
    assert declared, "no declarations found"
    assert os.path.exists(_lib.LIB_PATH), "build the library first (python -c 'import __graft_entry__ as g; g.build()')"
    L = ctypes.CDLL(_lib.LIB_PATH)
    for name in sorted(declared):
        assert hasattr(L, name), "missing export " + name
    assert set(_lib.EXPORTS) <= declared
    assert _lib.lib().seqm_version() >= 100
    assert _lib.lib().seqm_atompar_bytes() == 28 * 8


def test_no_cpu_fallback():
    from pyseqm_b200.basics import Energy
    from pyseqm_b200.seqm_functions.constants import Constants
    from pyseqm_b200 import synth
    sp_, xyz = synth.formaldehyde_pair()
    par = dict(method='AM1', scf_eps=1e-6, scf_converger=[2, 0.0], sp2=[False], elements=[0, 1, 6, 8], learned=[],
               pair_outer_cutoff=1e10)
    with pytest.raises(RuntimeError):
        Energy(par)(Constants(), torch.as_tensor(xyz), torch.as_tensor(sp_))
