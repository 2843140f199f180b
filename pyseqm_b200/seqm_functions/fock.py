"""fock() with the reference signature (seqm/seqm_functions/fock.py:6-139).  The stand-alone function gets a
dense w and no atomic numbers, so every pair is run through the heavy-heavy kernel (zeros in w make that
exact); Energy/Force/XL-BOMD use the class-tiled integrals directly.  Forward only."""
import torch

from .. import _ops
from .._topology import PairTopology


class _DenseW:
    def __init__(self, w):
        P = w.shape[0]
        T = (P + 31) // 32
        buf = torch.zeros(T * 32, 100, dtype=torch.float64, device=w.device)
        buf[:P] = w.reshape(P, 100)
        self.xx = buf.reshape(T, 32, 100).transpose(1, 2).contiguous()
        self.hh = torch.zeros(1, dtype=torch.float64, device=w.device)
        self.xh = torch.zeros(1, dtype=torch.float64, device=w.device)
        self.c = _ops.WStore(self.hh.data_ptr(), self.xh.data_ptr(), self.xx.data_ptr())


def fock(nmol, molsize, P0, M, maskd, mask, idxi, idxj, w, gss, gpp, gsp, gp2, hsp):
    if w.requires_grad or M.requires_grad or P0.requires_grad:
        raise NotImplementedError("stand-alone fock() is forward only; use Energy/Force for gradients")
    Z = torch.full_like(maskd, 6)
    topo = PairTopology(nmol, molsize, Z, maskd, idxi, idxj, all_heavy=True)
    return _ops.fock(topo, _DenseW(w), P0, M, (gss, gpp, gsp, gp2, hsp))
