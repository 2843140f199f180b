"""SP2 purification (seqm/seqm_functions/SP2.py:3-72) on packed matrices, FP64 tensor-core products."""
import torch

from .. import _ops


def SP2(a, nocc, eps=1.0e-4, factor=2.0):
    """a: packed symmetric matrices (nmol, m, m) (zero padded like pack() returns them); -> 2 X."""
    if factor != 2.0:
        raise NotImplementedError("SP2: only the closed-shell factor 2.0")
    B, m, _ = a.shape
    # the kernel takes padded-layout input: view the packed matrix as m//4 "heavy" atoms followed by m%4
    # one-orbital atoms
    nh, r = m // 4, m % 4
    S = nh + r
    n4 = 4 * S
    idx = torch.arange(m, device=a.device)
    idx = torch.where(idx < 4 * nh, idx, 4 * nh + 4 * (idx - 4 * nh))
    F = torch.zeros(B, n4, n4, dtype=a.dtype, device=a.device)
    F[:, idx.unsqueeze(1), idx.unsqueeze(0)] = a
    i32 = torch.int32
    nheavy = torch.full((B,), nh, dtype=i32, device=a.device)
    nhydro = torch.full((B,), r, dtype=i32, device=a.device)
    D = _ops.sp2_density(F, nheavy, nhydro, nocc.to(i32).contiguous(), S, m, eps)
    return D[:, idx][:, :, idx].contiguous()
