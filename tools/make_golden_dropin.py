"""Build-container script: run the user scripts of tests/dropin/ against the UNMODIFIED reference (imported as `seqm`
from /root/reference, CPU) and store what they return under tests/golden_dropin/.  tests/test_gpu_dropin.py runs the
same files with `sys.modules['seqm'] = pyseqm_b200`."""
import importlib.util
import os
import sys
import warnings

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
from ref_loader import load_reference  # noqa: E402

warnings.filterwarnings('ignore')
OUT = os.path.join(ROOT, 'tests', 'golden_dropin')


def load_script(name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(ROOT, 'tests', 'dropin', name + '.py'))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def main():
    os.makedirs(OUT, exist_ok=True)
    load_reference()
    cpu = torch.device('cpu')
    runs = [('script_energy_terms', 'energy_terms', dict(check_degeneracy=False)),
            ('script_energy_terms', 'energy_terms_degen', dict(check_degeneracy=True)),
            ('script_force_batch', 'force_batch', dict())]
    for script, tag, kw in runs:
        res = load_script(script).main(cpu, **kw)
        np.savez_compressed(os.path.join(OUT, tag + '.npz'), **{k: v.numpy() for k, v in res.items()})
        print(tag, {k: tuple(v.shape) for k, v in res.items()}, 'notconverged', res['notconverged'].tolist(),
              'Etot', res['Etot'].tolist())


if __name__ == '__main__':
    main()
