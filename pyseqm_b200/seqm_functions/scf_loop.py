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
    charge = torch.zeros(nmol, molsize * 4, molsize, device=v.device, dtype=v.dtype)
    v2 = v ** 2
    norb = 4 * nHeavy + nHydro
    for i in range(nmol):
        no, nX, nH = int(norb[i]), int(nHeavy[i]), int(nHydro[i])
        charge[i, :no, :nX] = v2[i][:no, :(4 * nX)].reshape(no, 4, nX).sum(dim=1)
        charge[i, :no, nX:(nX + nH)] = v2[i][:no, (4 * nX):(4 * nX + nH)]
    return charge


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
