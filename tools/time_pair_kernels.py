"""Times the pair-pass kernels (pair_fwd, pair_grad, fock_pair) of one Force call on 256 C60 (config 3's batch) with the
library's own event marks, for each library given on the command line (variant builds of tools/build_variant.sh; no
argument = the product library), each in its own process.  Prints one line per library: ms per launch group (median of
5 calls) and a checksum of the energies and forces so that a variant that changes the numbers shows up at once.

  python tools/time_pair_kernels.py [tools/_build/libseqm_x.so ...]
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def child():
    import ctypes
    import statistics
    import torch
    sys.path.insert(0, ROOT)
    torch.set_default_dtype(torch.float64)
    from pyseqm_b200 import synth, _lib
    from pyseqm_b200.basics import Force
    from pyseqm_b200.seqm_functions.constants import Constants
    dev = torch.device('cuda:0')
    const = Constants().to(dev)
    nmol = int(os.environ.get('SEQM_TIME_NMOL', '256'))
    sp, xyz = synth.c60(nmol, seed=0)
    par = dict(method='PM3', scf_eps=1.0e-6, scf_converger=[2, 0.0], sp2=[True, 1.0e-5], elements=[0, 6], learned=[],
               pair_outer_cutoff=1.0e10)
    force = Force(par).to(dev)
    coords, species = torch.as_tensor(xyz).to(dev), torch.as_tensor(sp).to(dev)
    for _ in range(2):
        out = force(const, coords, species)
    torch.cuda.synchronize()
    L = _lib.lib()
    acc = {}
    for _ in range(5):
        L.seqm_profile_enable(1)
        out = force(const, coords, species)
        buf = ctypes.create_string_buffer(1 << 16)
        L.seqm_profile_report(buf, len(buf))
        L.seqm_profile_enable(0)
        for line in buf.value.decode().splitlines():
            nm, ms, cnt = line.rsplit(' ', 2)
            acc.setdefault(nm, []).append(float(ms) / max(int(cnt), 1))
    keys = ['pair_fwd', 'pair_grad', 'fock_pair', 'hcore_diag', 'pair_geometry_grad']
    txt = ' '.join('%s %.4f' % (k, statistics.median(acc[k])) for k in keys if k in acc)
    f, e = out[0], out[2]
    print('%s | %s | sumE %.9f sum|f| %.9f' % (os.environ.get('SEQM_B200_LIB', 'product'), txt, float(e.sum()),
                                                 float(f.abs().sum())), flush=True)


if __name__ == '__main__':
    if os.environ.get('SEQM_TIME_CHILD') == '1':
        child()
    else:
        libs = sys.argv[1:] or [None]
        for lib in libs:
            env = dict(os.environ, SEQM_TIME_CHILD='1')
            if lib:
                env['SEQM_B200_LIB'] = os.path.abspath(lib)
            subprocess.run([sys.executable, os.path.abspath(__file__)], env=env, check=False, timeout=300)
