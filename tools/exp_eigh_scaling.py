"""eigh_density time against the number of molecules (latency regime vs throughput regime), benzene-size matrices."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
torch.set_default_dtype(torch.float64)
from pyseqm_b200 import _ops
dev = torch.device('cuda:0')
S, m = 12, 30
g = torch.Generator(device='cpu').manual_seed(1)
for nmol in [int(x) for x in os.environ.get("NMOLS", "37,148,296,592,1036,1184,2368,4736,9472").split(",")]:
    X = torch.randn(nmol, 48, 48, generator=g)
    X = X + X.transpose(1, 2)
    F = torch.zeros(nmol, 48, 48)
    idx = list(range(24)) + list(range(24, 48, 4))
    ii = torch.tensor(idx)
    F[:, ii[:, None], ii[None, :]] = X[:, :30, :30]
    F = F.to(dev)
    nH = torch.full((nmol,), 6, dtype=torch.int32, device=dev)
    nh = torch.full((nmol,), 6, dtype=torch.int32, device=dev)
    no = torch.full((nmol,), 15, dtype=torch.int32, device=dev)
    info = torch.zeros(nmol, dtype=torch.int32, device=dev)
    for _ in range(2):
        _ops.eigh_density(F, nH, nh, no, S, m, info=info)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(5):
        _ops.eigh_density(F, nH, nh, no, S, m)
    b.record()
    torch.cuda.synchronize()
    print('nmol %5d  %.3f ms  %.1f ns/molecule  sweeps min %d mean %.2f max %d' % (nmol, a.elapsed_time(b) / 5, a.elapsed_time(b) / 5 * 1e6 / nmol, int(info.min()), float(info.float().mean()), int(info.max())))
