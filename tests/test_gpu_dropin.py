"""The drop-in claim, tested as stated: user scripts written against the REFERENCE's import names (tests/dropin/*.py:
`import seqm`, `from seqm.basics import Energy`, module switches `seqm.seqm_functions.diag.CHECK_DEGENERACY`, the
`charges=` keyword, a finite `pair_outer_cutoff`, `Hf_flag`, `eig`) run UNCHANGED with `seqm` resolving to pyseqm_b200
on the GPU, and their results are compared with what the very same files returned when `seqm` was the unmodified
reference on the CPU (tests/golden_dropin/*.npz, tools/make_golden_dropin.py)."""
import importlib.util
import os
import sys

import numpy as np
import pytest
import torch

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _script(name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(ROOT, 'tests', 'dropin', name + '.py'))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.fixture()
def seqm_alias():
    import pyseqm_b200
    saved = {k: v for k, v in sys.modules.items() if k == 'seqm' or k.startswith('seqm.')}
    pyseqm_b200.install_as_seqm()
    yield
    for k in [k for k in sys.modules if k == 'seqm' or k.startswith('seqm.')]:
        del sys.modules[k]
    sys.modules.update(saved)


def _golden(tag):
    return np.load(os.path.join(ROOT, 'tests', 'golden_dropin', tag + '.npz'))


@pytest.mark.parametrize('degen', [False, True])
def test_energy_terms_script(seqm_alias, degen):
    """the reference's tests/test.py configuration: H2, C2, N2, O2 (PM3, adaptive mixing, charges kwarg, eig=True);
    C2 and O2 have a degenerate HOMO: with CHECK_DEGENERACY the occupation is smeared like diag.py:79-98"""
    res = _script('script_energy_terms').main(torch.device('cuda:0'), check_degeneracy=degen)
    g = _golden('energy_terms_degen' if degen else 'energy_terms')
    assert np.array_equal(res['notconverged'].numpy(), g['notconverged'])
    # C2 and O2 have a HOMO that is degenerate BY SYMMETRY.  Without the switch both codes fill it in eigenvalue order
    # and agree.  With the switch, whether the reference's 1e-14 test (diag.py:86) fires in a given SCF iteration
    # depends on the rounding noise LAPACK leaves between the two eigenvalues: its stored O2 result is a symmetry-broken
    # state (populations 1.00098 / 1.99902), while a Jacobi solver that returns the pair equal to the last bit smears
    # in every iteration and converges to the symmetric state, 5 eV higher.  Neither is reproducible from the other, so
    # with the switch on the strict comparison covers H2 and N2; the smearing arithmetic is pinned by the exact test below.
    keep = np.array([True, not degen, True, not degen])
    for k in ('Etot_m_Eiso', 'Etot', 'Eelec'):
        assert np.abs(res[k].numpy() - g[k])[keep].max() < 1e-6, k
    for k in ('Enuc', 'Eiso', 'EnucAB'):
        assert np.abs(res[k].numpy() - g[k]).max() < 1e-9, k
    assert res['e'].shape == g['e'].shape and res['charge'].shape == g['charge'].shape and res['P'].shape == g['P'].shape
    assert np.abs(res['e'].numpy() - g['e'])[keep].max() < 1e-5
    assert np.abs(np.trace(res['P'].numpy(), axis1=1, axis2=2) - np.trace(g['P'], axis1=1, axis2=2)).max() < 1e-8
    assert np.isfinite(res['Etot'].numpy()).all()


def test_check_degeneracy_smearing_exact(seqm_alias):
    """diag.CHECK_DEGENERACY on a matrix with an EXACTLY degenerate HOMO / LUMO pair: occupations (2, 1, 1, 0) with the
    switch (construct_P, diag.py:79-98: the degenerate set shares the remaining electrons), an integer occupation without"""
    import seqm
    from seqm.seqm_functions.diag import sym_eig_trunc
    dev = torch.device('cuda:0')
    F = torch.diag(torch.tensor([-3.0, -1.0, -1.0, 2.0], dtype=torch.float64, device=dev)).unsqueeze(0).repeat(2, 1, 1)
    one = torch.ones(2, dtype=torch.int64, device=dev)
    try:
        seqm.seqm_functions.diag.CHECK_DEGENERACY = True
        e, P, v = sym_eig_trunc(F, one, 0 * one, 2 * one)
        assert torch.allclose(P.diagonal(dim1=1, dim2=2).cpu(), torch.tensor([[2.0, 1.0, 1.0, 0.0]] * 2, dtype=torch.float64))
        assert torch.allclose(e.cpu(), torch.tensor([[-3.0, -1.0, -1.0, 2.0]] * 2, dtype=torch.float64))
        e, P, v = sym_eig_trunc(F, one, 0 * one, 3 * one)      # HOMO not degenerate with anything above it
        assert torch.allclose(P.diagonal(dim1=1, dim2=2).cpu(), torch.tensor([[2.0, 2.0, 2.0, 0.0]] * 2, dtype=torch.float64))
        seqm.seqm_functions.diag.CHECK_DEGENERACY = False
        e, P, v = sym_eig_trunc(F, one, 0 * one, 2 * one)
        d = P.diagonal(dim1=1, dim2=2).cpu()
        assert float(d[0, 0]) == 2.0 and sorted(d[0, 1:3].tolist()) == [0.0, 2.0] and float(d[0, 3]) == 0.0
    finally:
        seqm.seqm_functions.diag.CHECK_DEGENERACY = False


def test_force_batch_script(seqm_alias):
    res = _script('script_force_batch').main(torch.device('cuda:0'))
    g = _golden('force_batch')
    assert np.array_equal(res['notconverged'].numpy(), g['notconverged'])
    for k in ('Etot', 'Hf', 'Eelec'):
        assert np.abs(res[k].numpy() - g[k]).max() < 1e-6, k
    for k in ('Enuc', 'Eiso', 'EnucAB'):
        assert np.abs(res[k].numpy() - g[k]).max() < 1e-9, k
    assert np.abs(res['force'].numpy() - g['force']).max() < 1e-5
    assert np.abs(res['P'].numpy() - g['P']).max() < 1e-5
