"""The kernels' per-pair mathematics (pyseqm_b200/csrc/pair_math.cuh, compiled for the host by
tests/hostcheck) against the oracle: values of w, overlap, core repulsion, and their derivatives with
respect to (xij, rij) against the oracle's autograd.  Runs without a GPU."""
import numpy as np
import pytest
import torch

import seqm_oracle as orc
import hostcheck_util as hc


def _pairs_case(method, seed, axis_aligned=False):
    rng = np.random.RandomState(seed)
    Zs = [8, 7, 6, 1]
    pairs = [(zi, zj) for zi in Zs for zj in Zs if zi >= zj]
    ni = torch.tensor([p[0] for p in pairs])
    nj = torch.tensor([p[1] for p in pairs])
    n = len(pairs)
    v = rng.standard_normal((n, 3))
    if axis_aligned:
        v[:, :2] = 0.0
        v[::2, 2] = np.abs(v[::2, 2]) + 0.1
        v[1::2, 2] = -np.abs(v[1::2, 2]) - 0.1
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    xij = torch.tensor(v)
    rij = torch.tensor(rng.uniform(1.5, 9.0, n))
    tab = orc.load_params(method)
    return ni, nj, xij, rij, tab


def _atompars(tab, method, Zlist):
    Z = torch.tensor(Zlist)
    par = {k: v[Z] for k, v in tab.items()}
    dd, qq, r0, r1, r2 = orc.atom_multipoles(Z, par['zeta_s'], par['zeta_p'], par['g_ss'], par['g_pp'], par['g_p2'],
                                             par['h_sp'])
    ng = orc.N_GAUSS[method]
    out = []
    for a in range(len(Zlist)):
        K = par['K'][a].tolist() if ng else ()
        L = par['L'][a].tolist() if ng else ()
        M = par['M'][a].tolist() if ng else ()
        out.append(hc.atompar(float(par['zeta_s'][a]), float(par['zeta_p'][a]), float(par['beta_s'][a]),
                              float(par['beta_p'][a]), float(dd[a]), float(qq[a]), float(r0[a]), float(r1[a]),
                              float(r2[a]), float(par['alpha'][a]), float(orc.TORE[Zlist[a]]), K, L, M))
    return out, par, (dd, qq, r0, r1, r2)


@pytest.mark.parametrize('method', ['MNDO', 'AM1', 'PM3'])
@pytest.mark.parametrize('aligned', [False, True])
def test_pair_values_and_derivatives(method, aligned):
    ni, nj, xij, rij, tab = _pairs_case(method, 7, aligned)
    n = ni.shape[0]
    ng = orc.N_GAUSS[method]
    # oracle on a flat atom list: atoms 2p (first) and 2p+1 (second)
    Zflat = torch.stack((ni, nj), dim=1).reshape(-1)
    apars, par, (dd, qq, r0, r1, r2) = _atompars(tab, method, Zflat.tolist())
    idxi = torch.arange(n) * 2
    idxj = idxi + 1
    x = xij.clone().requires_grad_(True)
    r = rij.clone().requires_grad_(True)
    w, e1b, e2a = orc.two_center_integrals(ni, nj, x, r, dd, qq, r0, r1, r2, idxi, idxj)
    zeta = torch.stack((par['zeta_s'], par['zeta_p']), dim=1)
    di = orc.overlap(ni, nj, x, r, zeta[idxi], zeta[idxj])
    ps = dict(ni=ni, nj=nj, idxi=idxi, idxj=idxj, rij=r)
    en = orc.pair_nuclear_energy(ps, par, method, w[:, 0, 0])
    rng = np.random.RandomState(3)
    Gw = torch.tensor(rng.standard_normal((n, 10, 10)))
    Gd = torch.tensor(rng.standard_normal((n, 4, 4)))
    Ge = torch.tensor(rng.standard_normal(n))
    (w * Gw).sum().backward(retain_graph=True)
    gx_w, gr_w = x.grad.clone(), r.grad.clone()
    x.grad.zero_(); r.grad.zero_()
    (di * Gd).sum().backward(retain_graph=True)
    gx_d, gr_d = x.grad.clone(), r.grad.clone()
    x.grad.zero_(); r.grad.zero_()
    (en * Ge).sum().backward()
    gx_e, gr_e = x.grad.clone(), r.grad.clone()
    for p in range(n):
        a, b = apars[2 * p], apars[2 * p + 1]
        wv, dv, ev = hc.pair(int(ni[p]), int(nj[p]), xij[p].numpy(), float(rij[p]), a, b, ng)
        assert np.abs(wv - w[p].detach().numpy()).max() < 2e-13, (p, int(ni[p]), int(nj[p]))
        assert np.abs(dv - di[p].detach().numpy()).max() < 2e-13, (p, int(ni[p]), int(nj[p]))
        assert abs(ev - float(en[p].detach())) < 1e-11 * max(1.0, abs(float(en[p].detach())))
        wd, dd_, ed = hc.pair_dual(int(ni[p]), int(nj[p]), xij[p].numpy(), float(rij[p]), a, b, ng)
        assert np.abs(wd[..., 0] - wv).max() < 1e-14
        g = np.einsum('kl,klq->q', Gw[p].numpy(), wd[..., 1:])
        ref = np.concatenate([gx_w[p].numpy(), [float(gr_w[p])]])
        assert np.abs(g - ref).max() < 1e-10 * max(1.0, np.abs(ref).max()), ('w', p, g, ref)
        g = np.einsum('kl,klq->q', Gd[p].numpy(), dd_[..., 1:])
        ref = np.concatenate([gx_d[p].numpy(), [float(gr_d[p])]])
        assert np.abs(g - ref).max() < 1e-10 * max(1.0, np.abs(ref).max()), ('di', p, g, ref)
        g = float(Ge[p]) * ed[1:]
        ref = np.concatenate([gx_e[p].numpy(), [float(gr_e[p])]])
        assert np.abs(g - ref).max() < 1e-10 * max(1.0, np.abs(ref).max()), ('enuc', p, g, ref)


@pytest.mark.parametrize('aligned', [False, True])
def test_rotate_xx_adjoint_matches_forward_duals(aligned):
    """rotate_xx_adjoint (reverse mode, what pair_grad_xx_kernel runs) against the forward-mode dual derivative of the
    same rotation, heavy-heavy pairs, random upstream"""
    ni, nj, xij, rij, tab = _pairs_case('PM3', 11, aligned)
    Zflat = torch.stack((ni, nj), dim=1).reshape(-1)
    apars, par, _ = _atompars(tab, 'PM3', Zflat.tolist())
    rng = np.random.RandomState(5)
    checked = 0
    for p in range(ni.shape[0]):
        if int(ni[p]) <= 1 or int(nj[p]) <= 1:
            continue
        a, b = apars[2 * p], apars[2 * p + 1]
        G = rng.standard_normal((10, 10))
        wd, _, _ = hc.pair_dual(int(ni[p]), int(nj[p]), xij[p].numpy(), float(rij[p]), a, b, 2)
        ref = np.einsum('kl,klq->q', G, wd[..., 1:])
        got = hc.rotate_xx_adjoint(xij[p].numpy(), float(rij[p]), a, b, G)
        assert np.abs(got - ref).max() < 1e-12 * max(1.0, np.abs(ref).max()), (p, got, ref)
        checked += 1
    assert checked >= 6


@pytest.mark.parametrize('method', ['MNDO', 'PM3'])
@pytest.mark.parametrize('aligned', [False, True])
def test_overlap_xx_adjoint_matches_forward_duals(method, aligned):
    """overlap_xx_adjoint (radial integrals with one tangent, angular factors with three: what pair_grad_xx_kernel
    runs) against the four-tangent dual derivative of overlap_block, heavy-heavy pairs, random upstream; short and long
    distances (both branches of the auxiliary B integrals)"""
    ni, nj, xij, rij, tab = _pairs_case(method, 13, aligned)
    Zflat = torch.stack((ni, nj), dim=1).reshape(-1)
    apars, par, _ = _atompars(tab, method, Zflat.tolist())
    rng = np.random.RandomState(9)
    checked = 0
    for p in range(ni.shape[0]):
        if int(ni[p]) <= 1 or int(nj[p]) <= 1:
            continue
        a, b = apars[2 * p], apars[2 * p + 1]
        for r in (float(rij[p]), 0.9, 2.3, 12.0):
            F = rng.standard_normal((4, 4))
            _, dd, _ = hc.pair_dual(int(ni[p]), int(nj[p]), xij[p].numpy(), r, a, b, 0)
            ref = np.einsum('mn,mnq->q', F, dd[..., 1:])
            got = hc.overlap_xx_adjoint(xij[p].numpy(), r, a, b, F)
            assert np.abs(got - ref).max() < 1e-12 * max(1.0, np.abs(ref).max()), (p, r, got, ref)
            checked += 1
    assert checked >= 24
