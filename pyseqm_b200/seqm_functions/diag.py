"""sym_eig_trunc (seqm/seqm_functions/diag.py:102-172) on the batched shared-memory Jacobi kernel."""
import torch

from .. import _ops

CHECK_DEGENERACY = False


def sym_eig_trunc(x, nheavyatom, nH, nocc, eig_only=False):
    """x: padded Fock matrices (nmol, 4*molsize, 4*molsize).
    -> (e, P, v): orbital energies zero padded to 4*molsize, density 2 C_occ C_occ^T (padded layout),
    eigenvectors v[b, :m, :m] (packed rows, ascending columns).  eig_only=True -> (e, v)."""
    i32 = torch.int32
    molsize = x.shape[1] // 4
    mmax = int((4 * nheavyatom + nH).max())
    apply_check_degeneracy(mmax, sp2=False)
    P, e, v = _ops.eigh_density(x, nheavyatom.to(i32).contiguous(), nH.to(i32).contiguous(),
                                nocc.to(i32).contiguous(), molsize, mmax, want_eig=True)
    v = v[:, :mmax, :mmax]
    if eig_only:
        return e, v
    return e, P, v


def apply_check_degeneracy(mmax, sp2=False):
    """hand the module switch CHECK_DEGENERACY (reference diag.py:7-10) to the library before a density solve"""
    from .._lib import lib
    if CHECK_DEGENERACY and not sp2 and int(mmax) > lib().seqm_eigh_check_degeneracy_max_m():
        raise NotImplementedError("CHECK_DEGENERACY occupation smearing is implemented for packed sizes <= %d (warp "
                                  "eigensolver); this batch has %d" % (lib().seqm_eigh_check_degeneracy_max_m(), mmax))
    lib().seqm_eigh_check_degeneracy(1 if CHECK_DEGENERACY else 0)


def sym_eig_trunc1(*args, **kwargs):
    raise NotImplementedError("sym_eig_trunc1 (differentiable eigensolver for scf_backward>=1) is outside the "
                              "B200 hot path")
