"""Build-container helper: import the UNMODIFIED reference from /root/reference as `seqm`.
Used only by tools/make_golden.py and ad-hoc checks; nothing under tests/ or the product imports it."""
import sys
import warnings

REF_ROOT = '/root/reference'


def load_reference():
    warnings.filterwarnings('ignore', category=SyntaxWarning)
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    for k in [k for k in sys.modules if k == 'seqm' or k.startswith('seqm.')]:
        del sys.modules[k]
    import seqm  # noqa
    import seqm.basics, seqm.seqm_functions.hcore, seqm.seqm_functions.scf_loop  # noqa
    assert seqm.__file__.startswith(REF_ROOT), seqm.__file__
    return seqm
