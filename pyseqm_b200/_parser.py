"""Device-side Parser (csrc/parser_kernels.cu): the 16 tensors of reference basics.py:43-118 and the kernels' pair
topology from ONE pass over the atom slots, with the outer cutoff applied on the device.  Host synchronisations: one
(the sizes of the outputs and the validity flags travel in a single small transfer)."""
import ctypes

import torch

from . import _ops
from ._lib import Topology, check, lib
from ._topology import PairTopology

_p = _ops._p


def device_parse(const, species, coordinates, outercutoff, charges=None):
    """-> (the reference's 16-tuple, PairTopology)"""
    dev = coordinates.device
    B, S = species.shape
    sp = species.contiguous()
    R = coordinates.detach().contiguous()
    n = B * S
    i64, i32 = torch.int64, torch.int32
    lcf = float(const.length_conversion_factor)
    cnt = torch.empty(5, n, dtype=i32, device=dev)
    st = _ops._stream()
    check(lib().seqm_parser_count(B, S, _p(sp), _p(R), float(outercutoff), lcf, _p(cnt), st), 'seqm_parser_count')
    # exclusive scans (first-atom pairs per class, second-atom pairs, real atoms)
    inc = torch.cumsum(cnt.to(i64), dim=1)
    exc = inc - cnt
    off_cls, off_second, off_real = exc[0:3], exc[3], exc[4]
    # the pairs of a slot are written in b order whatever their class: their offset is the scan of the TOTAL per slot
    tot = cnt[0:3].sum(dim=0).to(i64)
    off_pair = torch.cumsum(tot, 0) - tot
    # per-molecule quantities
    nHeavy = torch.sum(species > 1, dim=1)
    nHydro = torch.sum(species == 1, dim=1)
    n_charge = torch.sum(const.tore[species], dim=1).reshape(-1).type(i64)
    if torch.is_tensor(charges):
        n_charge = n_charge - charges.reshape(-1).type(i64)
    nocc = torch.div(n_charge, 2, rounding_mode='floor')
    msz = 4 * nHeavy + nHydro
    zmax = species.max() if n else torch.zeros((), dtype=i64, device=dev)
    bad_el = ((species == 2)).any() | (zmax > 9)
    # heavy atoms must precede hydrogens inside a molecule (descending Z): a hydrogen followed by a heavy atom breaks it
    unsorted = ((species[:, :-1] == 1) & (species[:, 1:] > 1)).any() if S > 1 else torch.zeros((), dtype=torch.bool, device=dev)
    pack = torch.stack([inc[4, -1] if n else zmax * 0, inc[0, -1] if n else zmax * 0, inc[1, -1] if n else zmax * 0,
                        inc[2, -1] if n else zmax * 0, msz.max() if B else zmax * 0, msz.min() if B else zmax * 0,
                        ((n_charge % 2) == 1).any().to(i64), bad_el.to(i64), unsorted.to(i64)]).cpu().tolist()   # the one sync
    A, n0, n1, n2, mmax, mmin, odd, bad, uns = [int(v) for v in pack]
    if odd:
        raise ValueError("Only closed shell system (with even number of electrons) are supported")
    if bad:
        raise ValueError("\nError from diat.py, overlap matrix\nSome elements are not supported yet")
    if uns:
        raise ValueError("atoms must be sorted by descending atomic number inside every molecule")
    P = n0 + n1 + n2
    flat = species.reshape(-1)
    real = flat > 0
    real_atoms = torch.nonzero(real).reshape(-1) if A != n else torch.arange(n, device=dev, dtype=i64)
    Z = flat[real_atoms]
    mol_of = torch.div(real_atoms, S, rounding_mode='floor')
    slot_of = real_atoms - mol_of * S
    maskd = mol_of * (S * S) + slot_of * (S + 1)
    atom_molid = mol_of
    e64 = lambda *shape: torch.empty(*shape, dtype=i64, device=dev)
    e32 = lambda *shape: torch.empty(*shape, dtype=i32, device=dev)
    mask, pair_molid, ni, nj, idxi, idxj = (e64(P) for _ in range(6))
    xij = torch.empty(P, 3, dtype=torch.float64, device=dev)
    rij = torch.empty(P, dtype=torch.float64, device=dev)
    idxi32, idxj32, pair_cls, pair_slot, perm_j = (e32(P) for _ in range(5))
    cls_pairs = [e32(max(k, 0)) for k in (n0, n1, n2)]
    scratch = e32(n * S)
    check(lib().seqm_parser_fill(B, S, _p(sp), _p(R), float(outercutoff), lcf, _p(off_pair.contiguous()),
                                 _p(off_cls[0].contiguous()), _p(off_cls[1].contiguous()), _p(off_cls[2].contiguous()),
                                 _p(off_second.contiguous()), _p(off_real.contiguous()), _p(mask), _p(pair_molid), _p(ni),
                                 _p(nj), _p(idxi), _p(idxj), _p(xij), _p(rij), _p(idxi32), _p(idxj32), _p(pair_cls),
                                 _p(pair_slot), _p(cls_pairs[0]), _p(cls_pairs[1]), _p(cls_pairs[2]), _p(perm_j),
                                 _p(scratch), st), 'seqm_parser_fill')
    del scratch
    tail = lambda t, total: torch.cat([t[real_atoms], torch.full((1,), total, dtype=i64, device=dev)]).to(i32).contiguous()
    topo = PairTopology.from_device(B, S, A, P, Z.to(i32).contiguous(), mol_of.to(i32).contiguous(),
                                    slot_of.to(i32).contiguous(), idxi32, idxj32, pair_cls, pair_slot, cls_pairs,
                                    [n0, n1, n2], tail(off_pair, P), tail(off_second, P), perm_j, nHeavy, nHydro, nocc,
                                    mmax, mmin == mmax or B == 0)
    if coordinates.requires_grad:
        # the geometry must stay differentiable: recompute xij, rij through the kernel pair with the hand-written adjoint
        xij, rij = _ops.PairGeometry.apply(coordinates, topo, lcf)
    out = (B, S, nHeavy, nHydro, nocc, Z, maskd, atom_molid, mask, pair_molid, ni, nj, idxi, idxj, xij, rij)
    return out, topo
