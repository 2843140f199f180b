"""Host restatement (numpy, tests/dev/proto/onesided_ring.py) of the one-sided ring Jacobi of csrc/eigh_ring.cuh - shift,
recursive ring ordering, FP32 tangent, norm update that is exact for an inexact tangent, no rotation of noise-level
couplings with a large angle - against LAPACK: densities of Fock matrices with exactly degenerate orbitals (the case that
needs the last two rules) and of random symmetric matrices."""
import importlib.util
import os

import numpy as np
import pytest

from conftest import GOLDEN

HERE = os.path.dirname(os.path.abspath(__file__))
spec = importlib.util.spec_from_file_location('onesided_ring', os.path.join(HERE, 'dev', 'proto', 'onesided_ring.py'))
proto = importlib.util.module_from_spec(spec)
spec.loader.exec_module(proto)


def _check(A, nocc, tol_p):
    lam, V, real, sweeps = proto.jacobi_onesided(A, None, 16, tol=1e-16)
    P = proto.density(lam, V, real, nocc)
    e, U = np.linalg.eigh(A)
    Pref = 2 * U[:, :nocc] @ U[:, :nocc].T
    Vr = V[:, real]
    assert np.abs(P - Pref).max() < tol_p
    assert np.abs(np.sort(lam[real]) - e).max() < 5e-12
    assert np.abs(Vr.T @ Vr - np.eye(Vr.shape[1])).max() < 1e-13
    assert sweeps <= 10
    return sweeps


def test_degenerate_fock_matrices():
    import torch
    import seqm_oracle as orc
    g = np.load(os.path.join(GOLDEN, 'mixed12_MNDO_eigh_c2.npz'))      # holds CH4, NH3, ... : exactly degenerate orbitals
    Fp = orc.pack(torch.as_tensor(g['F']), torch.as_tensor(g['nHeavy']), torch.as_tensor(g['nHydro'])).numpy()
    for b in range(Fp.shape[0]):
        m = int(4 * g['nHeavy'][b] + g['nHydro'][b])
        _check(Fp[b][:m, :m], int(g['nocc'][b]), 1e-12)


@pytest.mark.parametrize('m', [5, 16, 30, 32])
def test_random_matrices(m):
    rng = np.random.RandomState(m)
    for trial in range(4):
        X = rng.standard_normal((m, m))
        X = X + X.T
        if trial == 0:
            X = X * 1e-3 + np.diag(rng.standard_normal(m) * 10.0)      # nearly diagonal
        e = np.linalg.eigvalsh(X)
        nocc = max(1, m // 2)
        gap = e[nocc] - e[nocc - 1] if nocc < m else 1.0
        _check(X, nocc, 1e-11 / min(1.0, gap))      # the density is as well conditioned as the gap allows
