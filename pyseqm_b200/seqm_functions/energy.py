"""Energy terms (mirror of seqm/seqm_functions/energy.py:4-118).  elec_energy and pair_nuclear_energy run
in CUDA kernels; the per-molecule sums are index_add plumbing on [nmol] vectors."""
import torch

from .constants import a0
from .. import _ops


def elec_energy_isolated_atom(const, Z, uss, upp, gss, gpp, gsp, gp2, hsp):
    return uss * const.ussc[Z] + upp * const.uppc[Z] + gss * const.gssc[Z] + gpp * const.gppc[Z] \
        + gsp * const.gspc[Z] + gp2 * const.gp2c[Z] + hsp * const.hspc[Z]


def elec_energy(P, F, Hcore):
    """0.5 * sum P (sym(Hcore) + F); Hcore is the upper-only padded matrix as hcore() returns it.
    Forward only (inside Energy/Force the differentiable path is _ops.PairEnergy)."""
    nmol, n4, _ = P.shape
    S = n4 // 4
    M = Hcore.reshape(nmol, S, 4, S, 4).transpose(2, 3).reshape(nmol * S * S, 4, 4).contiguous()
    return _ops.elec_energy(P, F, M, nmol, S)


def total_energy(nmol, pair_molid, EnucAB, Eelec):
    Enuc = torch.zeros((nmol,), dtype=EnucAB.dtype, device=EnucAB.device)
    Enuc = Enuc.index_add(0, pair_molid, EnucAB)
    return Eelec + Enuc, Enuc


def heat_formation(const, nmol, atom_molid, Z, Etot, Eiso, flag=True):
    Eiso_sum = torch.zeros_like(Etot).index_add(0, atom_molid, Eiso)
    if flag:
        eheat_sum = torch.zeros_like(Etot).index_add(0, atom_molid, const.eheat[Z])
        return Etot - Eiso_sum + eheat_sum, Eiso_sum
    return Etot - Eiso_sum, Eiso_sum
