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


def test_bucketed_batch_plan_covers_every_molecule_once():
    """host logic of pyseqm_b200.batching (config 4 sweeps): buckets partition the batch, keep every real atom, respect
    the size edges and the chunk cap"""
    import numpy as np
    import torch
    from pyseqm_b200 import synth
    from pyseqm_b200.batching import DEFAULT_EDGES, BucketedBatch, packed_sizes
    sp, _ = synth.mixed_organics(200, seed=3)
    species = torch.as_tensor(sp)
    bb = BucketedBatch(species, max_chunk=32)
    seen = torch.cat([idx for idx, _ in bb.buckets])
    assert sorted(seen.tolist()) == list(range(200))
    edges = [0] + list(DEFAULT_EDGES) + [1 << 30]
    for idx, bsp in bb.buckets:
        assert 0 < idx.numel() <= 32
        assert torch.equal((bsp > 0).sum(1), (species[idx] > 0).sum(1))          # no real atom trimmed
        assert int((bsp > 0).sum(1).max()) == bsp.shape[1]                        # padding trimmed to the bucket
        m = packed_sizes(bsp)
        k = int(np.searchsorted(edges, int(m.max()), side='left'))
        assert int(m.min()) > edges[k - 1] and int(m.max()) <= edges[k]
    sp2, xyz2 = synth.mixed_organics_sweep(100, seed=0, base=40)
    assert sp2.shape == (100, 30) and (sp2[40:80] == sp2[:40]).all() and not np.allclose(xyz2[40:80], xyz2[:40])


def test_new_entry_points_reject_bad_arguments_without_a_gpu():
    """argument checks of the sharded-SP2 / workspace entry points run before any CUDA call"""
    import ctypes
    from pyseqm_b200 import _lib
    L = _lib.lib()
    assert L.seqm_sp2_density_nccl(None, 0, 1, 4, 8, None, None, None, None, 1e-5, None, None, None, None, None) < 0
    assert b'ranks' in L.seqm_last_error()
    assert L.seqm_sp2_density_nccl(None, 2, 1, 4, 8, None, None, None, None, 1e-5, None, None, None, None, None) < 0
    assert b'bad argument' in L.seqm_last_error()
    h = ctypes.c_void_p()
    assert L.seqm_nccl_comm_create(None, 0, 2, ctypes.byref(h)) < 0
    assert L.seqm_nccl_comm_create(ctypes.create_string_buffer(128), 3, 2, ctypes.byref(h)) < 0
    assert L.seqm_nccl_unique_id(None) < 0
    # workspaces grow with the system and with the number of ranks' padding
    assert L.seqm_sp2_nccl_workspace(1, 1500, 8) > L.seqm_sp2_workspace(1, 1500) > 2 * 6000 * 6000 * 8
    assert L.seqm_elec_energy_workspace(1, 1500) == ((6000 * 6000 + 8191) // 8192) * 8
    assert L.seqm_elec_energy_workspace(1024, 12) == 1024 * 8


def test_pack_parameters_does_not_fill_the_callers_dict():
    """Pack_Parameters returns a new dict: the shared default `learned_parameters=dict()` of Energy / Force must stay
    empty (a filled default disables the cached-species path and is shared between BucketedBatch's worker threads)"""
    from pyseqm_b200.basics import Pack_Parameters
    pp = Pack_Parameters(dict(method='AM1', elements=[0, 1, 6], learned=[]))
    mine = {}
    out = pp(torch.tensor([6, 1, 1]), learned_params=mine)
    assert mine == {} and 'U_ss' in out and out['U_ss'].shape == (3,)
    out2 = pp(torch.tensor([1, 1]))
    assert out2['U_ss'].shape == (2,) and out['U_ss'].shape == (3,)
    assert pp.forward.__defaults__[0] == {}
