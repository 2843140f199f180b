"""hcore() with the reference signature (seqm/seqm_functions/hcore.py:6-167), running the pair-pass kernels.
Forward only; inside Energy/Force differentiation goes through _ops.PairEnergy."""
import torch

from .. import _ops
from .._topology import PairTopology


def hcore(const, nmol, molsize, maskd, mask, idxi, idxj, ni, nj, xij, rij, Z, zetas, zetap, uss, upp, gss, gpp,
          gp2, hsp, beta, Kbeta=None):
    """-> M[nmol*molsize**2, 4, 4] (upper blocks only), w[npairs, 10, 10]"""
    if Kbeta is not None:
        raise NotImplementedError("Kbeta (learned pair scaling) is outside the B200 hot path")
    if xij.requires_grad or rij.requires_grad:
        raise NotImplementedError("stand-alone hcore() is forward only; use Energy/Force for gradients")
    topo = PairTopology(nmol, molsize, Z, maskd, idxi, idxj)
    zero = torch.zeros_like(zetas)
    par = dict(zeta_s=zetas, zeta_p=zetap, beta_s=beta[:, 0].contiguous(), beta_p=beta[:, 1].contiguous(), g_ss=gss,
               g_pp=gpp, g_p2=gp2, h_sp=hsp, alpha=zero)
    atompar = _ops.atom_pack(topo, par, 'MNDO')
    w, M, _, _ = _ops.hcore(topo, atompar, xij, rij, uss, upp, 'MNDO', want_enuc=False)
    return M, _ops.w_expand(topo, w)
