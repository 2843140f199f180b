"""The cluster eigensolver (csrc/eigh_cluster_kernels.cu: 8-CTA cluster, one-sided block Jacobi over distributed
shared memory, 118 < m <= 256) against LAPACK (numpy.linalg.eigh on the host) on random symmetric matrices in the
padded layout of sym_eig_trunc (diag.py:102-172, pack.py:7-41)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
torch.set_default_dtype(torch.float64)


def _padded_index(nheavy, nhydro):
    return np.array(list(range(4 * nheavy)) + [4 * nheavy + 4 * k for k in range(nhydro)])


@pytest.mark.parametrize('nheavy,nhydro,nmol', [(30, 10, 3), (35, 20, 2), (40, 40, 2), (60, 0, 3), (64, 0, 2), (29, 3, 2)])
def test_cluster_eigh_matches_lapack(nheavy, nhydro, nmol):
    from pyseqm_b200 import _ops
    dev = torch.device('cuda:0')
    S = nheavy + nhydro + 1                       # one padding atom
    m = 4 * nheavy + nhydro
    assert 118 < m <= 256
    n4 = 4 * S
    rng = np.random.RandomState(m)
    idx = _padded_index(nheavy, nhydro)
    F = np.zeros((nmol, n4, n4))
    A = np.zeros((nmol, m, m))
    for b in range(nmol):
        a = rng.standard_normal((m, m))
        a = 0.5 * (a + a.T) * 3.0 + np.diag(np.linspace(-40.0, 25.0, m))      # eV-like spread, no exact degeneracy
        A[b] = a
        F[b][np.ix_(idx, idx)] = a
    nocc = m // 2 - 3
    i32 = lambda v: torch.full((nmol,), v, dtype=torch.int32, device=dev)
    P, e, v = _ops.eigh_density(torch.as_tensor(F).to(dev), i32(nheavy), i32(nhydro), i32(nocc), S, m, want_eig=True)
    torch.cuda.synchronize()
    for b in range(nmol):
        lam, C = np.linalg.eigh(A[b])
        assert np.abs(e[b, :m].cpu().numpy() - lam).max() < 1e-10
        assert float(e[b, m:].abs().max()) == 0.0
        Pref = 2.0 * C[:, :nocc] @ C[:, :nocc].T
        Pg = P[b].cpu().numpy()
        assert np.abs(Pg[np.ix_(idx, idx)] - Pref).max() < 1e-10
        mask = np.ones((n4, n4), dtype=bool)
        mask[np.ix_(idx, idx)] = False
        assert np.abs(Pg[mask]).max() == 0.0
        # eigenvectors: columns up to sign
        Vg = v[b, :m, :m].cpu().numpy()
        assert np.abs(np.abs(np.sum(Vg * C, axis=0)) - 1.0).max() < 1e-9


def test_cluster_eigh_mixed_sizes_and_active_mask():
    """one launch, molecules of different packed size (own m, common panel width) and an inactive molecule"""
    from pyseqm_b200 import _ops
    dev = torch.device('cuda:0')
    sizes = [(60, 0), (40, 25), (31, 2), (50, 10)]
    S = 66
    n4 = 4 * S
    rng = np.random.RandomState(7)
    F = np.zeros((len(sizes), n4, n4))
    ref = []
    for b, (nx, nh) in enumerate(sizes):
        m = 4 * nx + nh
        idx = _padded_index(nx, nh)
        a = rng.standard_normal((m, m))
        a = 0.5 * (a + a.T) * 2.0 + np.diag(np.linspace(-30.0, 20.0, m))
        F[b][np.ix_(idx, idx)] = a
        ref.append((idx, a, m))
    nheavy = torch.tensor([s[0] for s in sizes], dtype=torch.int32, device=dev)
    nhydro = torch.tensor([s[1] for s in sizes], dtype=torch.int32, device=dev)
    nocc = torch.tensor([(4 * s[0] + s[1]) // 2 for s in sizes], dtype=torch.int32, device=dev)
    active = torch.tensor([1, 1, 0, 1], dtype=torch.uint8, device=dev)
    P = _ops.eigh_density(torch.as_tensor(F).to(dev), nheavy, nhydro, nocc, S, 240, active=active)
    torch.cuda.synchronize()
    for b, (idx, a, m) in enumerate(ref):
        Pg = P[b].cpu().numpy()
        if b == 2:
            assert np.abs(Pg).max() == 0.0
            continue
        lam, C = np.linalg.eigh(a)
        no = m // 2
        assert np.abs(Pg[np.ix_(idx, idx)] - 2.0 * C[:, :no] @ C[:, :no].T).max() < 1e-10
