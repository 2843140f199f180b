import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, 'oracle')):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, 'tests', 'golden')
GOLDEN_BIG = os.path.join(ROOT, 'tests', 'golden_big')      # compact fixtures of the large cases (tools/make_golden_big.py)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def golden_names():
    return sorted(f[:-4] for f in os.listdir(GOLDEN) if f.endswith('.npz'))


def golden_big_names():
    return sorted(f[:-4] for f in os.listdir(GOLDEN_BIG) if f.endswith('.npz'))


def probe_vector(n):
    """the fixed vector the compact fixtures project P and F on (tools/make_golden_big.py)"""
    import numpy as np
    return 1.0 + 0.5 * np.sin(np.arange(n, dtype=np.float64) + 1.0)


def nondegenerate_molecules(e, nHeavy, nHydro, tol=1e-6):
    """mask of the molecules whose orbital energies are pairwise distinct"""
    import numpy as np
    ok = np.ones(e.shape[0], dtype=bool)
    for b in range(e.shape[0]):
        no = int(4 * nHeavy[b] + nHydro[b])
        ok[b] = bool((np.diff(np.sort(e[b, :no])) > tol).all())
    return ok


@pytest.fixture(scope='session')
def golden_dir():
    return GOLDEN
