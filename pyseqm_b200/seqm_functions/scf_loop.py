"""scf_loop with the reference signature (seqm/seqm_functions/scf_loop.py:671-806), backed by the
device-resident SCF driver (csrc/scf_kernels.cu).  Module knobs `debug` and `MAX_ITER` are honoured like
the reference's."""
import torch

from .. import _engine, _ops

debug = False
MAX_ITER = 1000
RAISE_ERROR_IF_SCF_FORWARD_FAILS = False


def _charges(v, nHeavy, nHydro, nmol, molsize):
    """orbital populations per atom (scf_loop.py:788-800), from the packed eigenvectors"""
    # literal index arithmetic of the reference, for all molecules at once (no host loop, no synchronisation):
    # charge[b, r, a] = sum_c v2[b, r, c nX + a] (a < nX: the reshape(norb, 4, nX).sum(1)), = v2[b, r, 4 nX + a - nX] for
    # the hydrogens, rows r < norb
    dev = v.device
    n = 4 * molsize
    mm = v.shape[-1]
    v2 = torch.zeros(nmol, n, n, device=dev, dtype=v.dtype)
    v2[:, :mm, :mm] = v ** 2
    a = torch.arange(molsize, device=dev).view(1, molsize, 1)
    c = torch.arange(4, device=dev).view(1, 1, 4)
    nX, nH = nHeavy.view(-1, 1, 1).long(), nHydro.view(-1, 1, 1).long()
    heavy, hyd = a < nX, (a >= nX) & (a < nX + nH)
    col = torch.where(heavy, c * nX + a, (4 * nX + a - nX).expand(-1, -1, 4)).clamp(0, n - 1)
    wgt = (heavy | (hyd & (c == 0))).to(v.dtype)
    idx = col.reshape(nmol, 1, 4 * molsize).expand(nmol, n, 4 * molsize)
    g = torch.gather(v2, 2, idx).reshape(nmol, n, molsize, 4) * wgt.unsqueeze(1)
    rows = (torch.arange(n, device=dev).view(1, n, 1) < (4 * nX + nH).view(-1, 1, 1)).to(v.dtype)
    return g.sum(dim=3) * rows


def scf_loop(const, molsize, nHeavy, nHydro, nOccMO, maskd, mask, atom_molid, pair_molid, idxi, idxj, ni, nj, xij,
             rij, Z, zetas, zetap, uss, upp, gss, gsp, gpp, gp2, hsp, beta, Kbeta=None, eps=1.0e-4, P=None,
             sp2=[False], scf_converger=[1], eig=False, scf_backward=0, scf_backward_eps=1.0e-2):
    """-> F, e, P, Hcore, w, charge, notconverged  (e, charge are None unless eig)"""
    if scf_backward != 0:
        raise NotImplementedError("scf_backward=1/2 (gradients through the SCF iterations) is outside the B200 "
                                  "hot path; forces use scf_backward=0 (Hellmann-Feynman) like the reference default")
    if Kbeta is not None:
        raise NotImplementedError("Kbeta is outside the B200 hot path")
    nmol = nOccMO.shape[0]
    zero = torch.zeros_like(zetas)
    par = dict(zeta_s=zetas, zeta_p=zetap, beta_s=beta[:, 0].contiguous(), beta_p=beta[:, 1].contiguous(), U_ss=uss,
               U_pp=upp, g_ss=gss, g_sp=gsp, g_pp=gpp, g_p2=gp2, h_sp=hsp, alpha=zero)
    st = _engine.build(const, nmol, molsize, nHeavy, nHydro, nOccMO, Z, maskd, idxi, idxj, xij, rij, par, 'MNDO',
                       run_scf=True, P0=P, eps=float(eps), sp2=sp2, scf_converger=scf_converger, max_iter=MAX_ITER,
                       debug=debug)
    if RAISE_ERROR_IF_SCF_FORWARD_FAILS and bool(st['notconverged'].any()):
        raise ValueError("SCF for some the molecules in the batch doesn't converge")
    Hcore = _engine.hcore_matrix(st['M'], nmol, molsize)
    w = _ops.w_expand(st['topo'], st['w'])
    e = charge = None
    if eig:
        from .diag import sym_eig_trunc
        e, v = sym_eig_trunc(st['F'], nHeavy, nHydro, nOccMO, eig_only=True)
        charge = _charges(v, nHeavy, nHydro, nmol, molsize)
    return st['F'], e, st['P'], Hcore, w, charge, st['notconverged']
