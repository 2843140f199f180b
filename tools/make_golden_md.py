"""Golden trajectories for the MD drivers from the UNMODIFIED reference: BOMD (Molecular_Dynamics_Basic) and
XL-BOMD (XL_BOMD) on the two H2CO-class molecules of tests/test8.py / test9.py, with fixed velocities.

    python tools/make_golden_md.py
"""
import contextlib
import io
import os
import sys
import warnings

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
from ref_loader import load_reference  # noqa: E402
from pyseqm_b200 import synth  # noqa: E402

torch.set_default_dtype(torch.float64)
warnings.filterwarnings('ignore')
OUT = os.path.join(ROOT, 'tests', 'golden_md')


def main():
    os.makedirs(OUT, exist_ok=True)
    seqm = load_reference()
    from seqm.seqm_functions.constants import Constants
    from seqm.MolecularDynamics import Molecular_Dynamics_Basic
    from seqm.XLBOMD import XL_BOMD
    sp_, xyz = synth.formaldehyde_pair()
    sp_, xyz = sp_[:1].repeat(2, 0), xyz[:1].repeat(2, 0)
    xyz[1] += 0.03 * np.random.RandomState(5).standard_normal(xyz[1].shape)
    vel = 0.01 * np.random.RandomState(7).standard_normal(xyz.shape)
    only = sys.argv[1:]
    # config 3 at its own molecule size: 2 C60, PM3, SP2 1e-5, XL-BOMD k = 5 (BASELINE.json configs[2])
    spc, xyzc = synth.c60(2, seed=8)
    velc = 0.004 * np.random.RandomState(9).standard_normal(xyzc.shape)
    cases = (('bomd_eigh', Molecular_Dynamics_Basic, [False], {}, 4, 'AM1', sp_, xyz, vel, 1e-8),
             ('xlbomd_sp2', XL_BOMD, [True, 1e-7], dict(k=5), 6, 'AM1', sp_, xyz, vel, 1e-8),
             ('xlbomd_eigh', XL_BOMD, [False], dict(k=5), 6, 'AM1', sp_, xyz, vel, 1e-8),
             ('xlbomd_c60_sp2', XL_BOMD, [True, 1e-5], dict(k=5), 10, 'PM3', spc, xyzc, velc, 1e-6))
    for tag, cls, sp2, kw, steps, method, sp_, xyz, vel, eps in cases:
        if only and tag not in only:
            continue
        par = dict(method=method, scf_eps=eps, scf_converger=[2, 0.0], sp2=sp2, elements=[0, 1, 6, 8], learned=[],
                   parameter_file_dir=os.path.join(seqm.__path__[0], 'params') + '/', pair_outer_cutoff=1e10)
        const = Constants()
        out = {'molid': [0, 1], 'thermo': 1, 'dump': 10 ** 9, 'prefix': '/tmp/golden_md'}
        md = cls(par, timestep=1.0, output=out, **kw)
        c = torch.as_tensor(xyz).clone()
        v = torch.as_tensor(vel).clone()
        buf = io.StringIO()
        with contextlib.redirect_stdout(buf):
            res = md.run(const, steps, c, v, torch.as_tensor(sp_))
        thermo = []
        for line in buf.getvalue().splitlines():
            t = line.split()
            if len(t) == 15 and t[0].isdigit():
                thermo.append([float(x) for x in t[1:]])
        np.savez_compressed(os.path.join(OUT, tag + '.npz'), species=sp_, coords0=xyz, vel0=vel, steps=steps,
                            sp2=np.array([float(x) for x in sp2]), k=kw.get('k', 0), method=method, eps=eps,
                            coords=res[0].detach().numpy(), vel=res[1].detach().numpy(), thermo=np.array(thermo))
        print(tag, 'steps', steps, 'thermo rows', len(thermo), 'Epot last', thermo[-1][2], thermo[-1][9])


if __name__ == '__main__':
    main()
