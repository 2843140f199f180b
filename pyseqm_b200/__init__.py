"""pyseqm_b200 — B200-native (sm_100a) implementation of PYSEQM's batched SCF hot path.

Mirrors the reference's `seqm` package layout for that path (`basics`, `seqm_functions.*`,
`MolecularDynamics`, `XLBOMD`); every numerical op runs in hand-written CUDA kernels reached through
the C-ABI library `libseqm_b200.so` (see include/seqm_b200.h).  There is no CPU fallback: using an op
without the library, or with CPU tensors, raises.
"""
__version__ = "0.1.0"


def install_as_seqm():
    """Register this package under the reference's import name: after the call `import seqm`,
    `from seqm.basics import Force`, `seqm.seqm_functions.diag.CHECK_DEGENERACY = True` ... resolve to pyseqm_b200
    (tests/test_gpu_dropin.py runs reference-style user scripts this way).  Returns the package."""
    import importlib
    import sys
    pkg = sys.modules[__name__]
    names = ['basics', 'MolecularDynamics', 'XLBOMD', 'seqm_functions', 'seqm_functions.constants',
             'seqm_functions.parameters', 'seqm_functions.hcore', 'seqm_functions.fock', 'seqm_functions.diag',
             'seqm_functions.SP2', 'seqm_functions.pack', 'seqm_functions.energy', 'seqm_functions.scf_loop']
    sys.modules['seqm'] = pkg
    for n in names:
        sys.modules['seqm.' + n] = importlib.import_module(__name__ + '.' + n)
    return pkg
