"""Step-by-step check of the two-warp kernels (run every stage under `timeout`):
  stage eigh : stand-alone two-warp Jacobi on random batches against torch.linalg.eigh
  stage scf  : Force on one golden case with the two-warp SCF kernel and a small max_iter"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..'))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..', 'oracle'))
torch.set_default_dtype(torch.float64)
dev = torch.device('cuda:0')


def stage_eigh():
    from pyseqm_b200 import _lib, _ops
    L = _lib.lib()
    for mmax in (8, 12, 20, 30, 32):
        rng = np.random.RandomState(mmax)
        B = 64
        a, h = mmax // 4, mmax % 4
        S = a + h
        n4 = 4 * S
        F = torch.zeros(B, n4, n4)
        idx = [4 * i + k for i in range(a) for k in range(4)] + [4 * (a + i) for i in range(h)]
        ii = torch.tensor(idx)
        for b in range(B):
            X = torch.as_tensor(rng.standard_normal((mmax, mmax)))
            X = X + X.T
            F[b][ii[:, None], ii[None, :]] = X
        nH = torch.full((B,), a, dtype=torch.int32, device=dev)
        nh = torch.full((B,), h, dtype=torch.int32, device=dev)
        no = torch.full((B,), mmax // 2, dtype=torch.int32, device=dev)
        Fd = F.to(dev)
        res = {}
        for warps in (1, 2):
            L.seqm_scf_warps(warps)
            info = torch.zeros(B, dtype=torch.int32, device=dev)
            t0 = time.time()
            P = _ops.eigh_density(Fd, nH, nh, no, S, mmax, info=info)
            torch.cuda.synchronize()
            res[warps] = (P.cpu(), info.cpu(), time.time() - t0)
        Fp = F[:, ii[:, None], ii[None, :]]
        e, v = torch.linalg.eigh(Fp)
        C = v[:, :, :mmax // 2]
        Pref = 2.0 * C @ C.transpose(1, 2)
        for warps in (1, 2):
            P, info, dt = res[warps]
            Pp = P[:, ii[:, None], ii[None, :]]
            print('eigh m=%d warps=%d  max|P-Pref|=%.3e  sweeps %d..%d  %.3fs' % (
                mmax, warps, float((Pp - Pref).abs().max()), int(info.min()), int(info.max()), dt), flush=True)


def stage_scf(name, max_iter):
    from pyseqm_b200 import _lib
    from pyseqm_b200.basics import Force
    from pyseqm_b200.seqm_functions.constants import Constants
    from pyseqm_b200.seqm_functions import scf_loop
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..'))
    import test_gpu_parity as T
    gg = T._load(name)
    sp = T._params(gg)
    scf_loop.MAX_ITER = max_iter
    const = Constants().to(dev)
    for warps in (1, 2):
        _lib.lib().seqm_scf_warps(warps)
        species, coords = T._t(gg['species'], dev), T._t(gg['coords'], dev)
        fm = Force(sp).to(dev)
        t0 = time.time()
        out = fm(const, coords, species)
        torch.cuda.synchronize()
        st = fm.energy.last_state
        print('scf %s warps=%d iterations=%s notconv=%d  dE=%.3e dP=%.3e dF=%.3e  %.3fs' % (
            name, warps, st['iterations'], int(out[-1].sum()),
            float((out[2].detach().cpu() - torch.as_tensor(gg['Etot'])).abs().max()),
            float((out[1].detach().cpu() - torch.as_tensor(gg['P'])).abs().max()),
            float((out[0].detach().cpu() - torch.as_tensor(gg['force'])).abs().max()), time.time() - t0), flush=True)


if __name__ == '__main__':
    if sys.argv[1] == 'eigh':
        stage_eigh()
    else:
        stage_scf(sys.argv[2], int(sys.argv[3]))
