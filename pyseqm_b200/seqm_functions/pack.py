"""pack / unpack (seqm/seqm_functions/pack.py:7-41) as batched gathers: no per-molecule Python loop.
Inside the kernels the same index map is folded into the loads/stores; these functions exist for callers
that use them directly."""
import torch


def _index(nHeavy, nHydro, n4):
    B = nHeavy.shape[0]
    norb = 4 * nHeavy + nHydro
    mmax = int(norb.max())
    q = torch.arange(mmax, device=nHeavy.device).unsqueeze(0).expand(B, mmax)
    nh4 = (4 * nHeavy).unsqueeze(1)
    idx = torch.where(q < nh4, q, nh4 + 4 * (q - nh4))
    ok = q < norb.unsqueeze(1)
    return idx, ok, mmax


def pack(x, nHeavy, nHydro):
    """(nmol, 4*molsize, 4*molsize) -> (nmol, mmax, mmax), zero padded"""
    B, n4, _ = x.shape
    idx, ok, m = _index(nHeavy, nHydro, n4)
    g = torch.where(ok, idx, torch.zeros_like(idx))
    y = x.gather(1, g.unsqueeze(2).expand(B, m, n4)).gather(2, g.unsqueeze(1).expand(B, m, m))
    return y * (ok.unsqueeze(2) & ok.unsqueeze(1)).to(x.dtype)


def unpack(x, nHeavy, nHydro, size):
    """(nmol, mmax, mmax) -> (nmol, size, size)"""
    B, m, _ = x.shape
    idx, ok, mm = _index(nHeavy, nHydro, size)
    idx, ok = idx[:, :m], ok[:, :m]
    if mm < m:
        pad = m - mm
        idx = torch.cat((idx, torch.zeros(B, pad, dtype=idx.dtype, device=idx.device)), 1)
        ok = torch.cat((ok, torch.zeros(B, pad, dtype=torch.bool, device=ok.device)), 1)
    g = torch.where(ok, idx, torch.full_like(idx, size))
    out = torch.zeros(B, size + 1, size + 1, dtype=x.dtype, device=x.device)
    b = torch.arange(B, device=x.device).reshape(B, 1, 1).expand(B, m, m)
    out[b, g.unsqueeze(2).expand(B, m, m), g.unsqueeze(1).expand(B, m, m)] = x
    return out[:, :size, :size].contiguous()
