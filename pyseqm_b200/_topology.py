"""Pair/atom topology handed to the kernels (seqm_topology_t): the Parser's int64 index tensors
(reference seqm/basics.py:43-118) narrowed to int32, pairs grouped by class (HH / XH / XX) and the CSR
tables the fixed-order per-atom reductions use.  Built with torch index ops on the device, once per pair
list."""
import ctypes

import torch

from ._lib import Topology

CLS_HH, CLS_XH, CLS_XX = 0, 1, 2


class PairTopology:
    def __init__(self, nmol, molsize, Z, maskd, idxi, idxj, nHeavy=None, nHydro=None, nocc=None, all_heavy=False):
        dev = Z.device
        if dev.type != 'cuda':
            raise RuntimeError("pyseqm_b200 runs on CUDA tensors only (there is no CPU path)")
        i32 = torch.int32
        S = int(molsize)
        self.nmol, self.molsize = int(nmol), S
        self.natoms, self.npairs = int(Z.shape[0]), int(idxi.shape[0])
        if int(Z.max()) > 9 or bool(((Z != 1) & (Z < 3)).any()):
            # reference diat_overlap.py:65-72: only valence shells n = 1, 2 are implemented
            raise ValueError("\nError from diat.py, overlap matrix\nSome elements are not supported yet")
        self.Z = Z.to(i32).contiguous()
        mol = torch.div(maskd, S * S, rounding_mode='floor')
        self.atom_mol = mol.to(i32).contiguous()
        self.atom_slot = torch.div(maskd - mol * S * S, S + 1, rounding_mode='floor').to(i32).contiguous()
        self.idxi = idxi.to(i32).contiguous()
        self.idxj = idxj.to(i32).contiguous()
        if all_heavy:       # generic topology for the stand-alone fock(): every pair treated as heavy-heavy
            Z = torch.full_like(Z, 6)
            self.Z = Z.to(i32).contiguous()
        hi = (Z[idxi] > 1).to(i32)
        hj = (Z[idxj] > 1).to(i32)
        if bool((hj > hi).any()):
            raise ValueError("atoms must be sorted by descending atomic number inside every molecule")
        cls = hi + hj
        self.pair_cls = cls.contiguous()
        self.cls_pairs = []
        slot = torch.zeros(self.npairs, dtype=i32, device=dev)
        for c in range(3):
            ids = torch.nonzero(cls == c).reshape(-1).to(i32).contiguous()
            self.cls_pairs.append(ids)
            slot[ids.long()] = torch.arange(ids.shape[0], dtype=i32, device=dev)
        self.pair_slot = slot
        self.ncls = [int(t.shape[0]) for t in self.cls_pairs]
        ar = torch.arange(self.natoms + 1, device=dev, dtype=idxi.dtype)
        self.start_i = torch.searchsorted(idxi.contiguous(), ar).to(i32).contiguous()
        sj, perm = torch.sort(idxj, stable=True)
        self.perm_j = perm.to(i32).contiguous()
        self.start_j = torch.searchsorted(sj.contiguous(), ar).to(i32).contiguous()
        if nHeavy is not None:
            self.nheavy = nHeavy.to(i32).contiguous()
            self.nhydro = nHydro.to(i32).contiguous()
            self.nocc = nocc.to(i32).contiguous()
            msz = 4 * nHeavy + nHydro
            self.mmax = int(msz.max()) if self.nmol else 0
            self.uniform = bool(self.nmol == 0 or int(msz.min()) == self.mmax)      # every molecule has the same packed size
        t = Topology()
        t.nmol, t.molsize, t.natoms, t.npairs = self.nmol, S, self.natoms, self.npairs
        t.Z, t.atom_mol, t.atom_slot = self.Z.data_ptr(), self.atom_mol.data_ptr(), self.atom_slot.data_ptr()
        t.idxi, t.idxj = self.idxi.data_ptr(), self.idxj.data_ptr()
        t.pair_cls, t.pair_slot = self.pair_cls.data_ptr(), self.pair_slot.data_ptr()
        for c in range(3):
            t.cls_pairs[c] = self.cls_pairs[c].data_ptr()
            t.ncls[c] = self.ncls[c]
        t.start_i, t.start_j, t.perm_j = self.start_i.data_ptr(), self.start_j.data_ptr(), self.perm_j.data_ptr()
        self.c = t
        self.device = dev

    @classmethod
    def from_device(cls, nmol, molsize, natoms, npairs, Z, atom_mol, atom_slot, idxi, idxj, pair_cls, pair_slot,
                    cls_pairs, ncls, start_i, start_j, perm_j, nHeavy, nHydro, nocc, mmax, uniform):
        """topology whose tables were produced on the device by the Parser kernels (_parser.device_parse): int32 tensors
        taken as they are, no index arithmetic and no synchronisation here"""
        self = cls.__new__(cls)
        i32 = torch.int32
        self.nmol, self.molsize, self.natoms, self.npairs = int(nmol), int(molsize), int(natoms), int(npairs)
        self.Z, self.atom_mol, self.atom_slot, self.idxi, self.idxj = Z, atom_mol, atom_slot, idxi, idxj
        self.pair_cls, self.pair_slot, self.cls_pairs, self.ncls = pair_cls, pair_slot, list(cls_pairs), [int(k) for k in ncls]
        self.start_i, self.start_j, self.perm_j = start_i, start_j, perm_j
        self.nheavy = nHeavy.to(i32).contiguous()
        self.nhydro = nHydro.to(i32).contiguous()
        self.nocc = nocc.to(i32).contiguous()
        self.mmax, self.uniform = int(mmax), bool(uniform)
        self.device = Z.device
        self._fill_struct()
        return self

    def _fill_struct(self):
        t = Topology()
        t.nmol, t.molsize, t.natoms, t.npairs = self.nmol, self.molsize, self.natoms, self.npairs
        t.Z, t.atom_mol, t.atom_slot = self.Z.data_ptr(), self.atom_mol.data_ptr(), self.atom_slot.data_ptr()
        t.idxi, t.idxj = self.idxi.data_ptr(), self.idxj.data_ptr()
        t.pair_cls, t.pair_slot = self.pair_cls.data_ptr(), self.pair_slot.data_ptr()
        for c in range(3):
            t.cls_pairs[c] = self.cls_pairs[c].data_ptr()
            t.ncls[c] = self.ncls[c]
        t.start_i, t.start_j, t.perm_j = self.start_i.data_ptr(), self.start_j.data_ptr(), self.perm_j.data_ptr()
        self.c = t

    def ref(self):
        return ctypes.byref(self.c)

    def w_sizes(self):
        """element counts of the class-tiled integral arrays (wHH, wXH, wXX)"""
        t = lambda n, e: ((n + 31) // 32) * 32 * e
        return max(self.ncls[0], 1), max(t(self.ncls[1], 10), 1), max(t(self.ncls[2], 100), 1)
