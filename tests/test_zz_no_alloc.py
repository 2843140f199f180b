"""seqm_elec_energy_noalloc - the allocation-free variant of seqm_elec_energy whose partial sums live in a file-scope
`static __device__` array (include/seqm_b200.h; no counterpart in the reference)
- called through the C ABI and compared with the existing entry point seqm_elec_energy on the same
inputs, on two calls in a row (the static array is shared between calls and keeps the previous call's contents).

Collected last on purpose (file name).  The test builds its own inputs, so it passes
alone in a fresh process as well as after the other tests."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 1.0e-9      # eV: the tolerance the other GPU tests use for electronic energies


def _p(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


def _inputs(nmol, molsize, seed, dev):
    """random symmetric P, F [B, 4S, 4S] and block matrix M [B*S*S, 4, 4] (any numbers do: the op is a weighted sum)"""
    g = torch.Generator().manual_seed(seed)
    n4 = 4 * molsize
    sym = lambda: (lambda a: 0.5 * (a + a.transpose(1, 2)))(torch.randn(nmol, n4, n4, generator=g, dtype=torch.float64))
    P, F, D = sym().to(dev).contiguous(), (10.0 * sym()).to(dev).contiguous(), sym().to(dev).contiguous()
    M = (5.0 * torch.randn(nmol * molsize * molsize, 4, 4, generator=g, dtype=torch.float64)).to(dev).contiguous()
    return P, F, D, M


def _both(L, nmol, molsize, P, F, D, M, xl, active, stream):
    dev = P.device
    e_ref = torch.full((nmol,), -7.0, dtype=torch.float64, device=dev)
    e_new = torch.full((nmol,), -7.0, dtype=torch.float64, device=dev)
    nbytes = L.seqm_elec_energy_workspace(nmol, molsize)
    work = torch.empty(max(int(nbytes), 8), dtype=torch.uint8, device=dev)      # the EXISTING entry point's caller scratch
    rc = L.seqm_elec_energy(nmol, molsize, _p(P), _p(F), _p(M), _p(D if xl else None), int(xl), _p(active), _p(e_ref),
                            _p(work), stream)
    assert rc == 0, L.seqm_last_error()
    rc = L.seqm_elec_energy_noalloc(nmol, molsize, _p(P), _p(F), _p(M), _p(D if xl else None), int(xl), _p(active),
                                    _p(e_new), stream)
    assert rc == 0, L.seqm_last_error()
    torch.cuda.synchronize()
    return e_ref.cpu().numpy(), e_new.cpu().numpy()


def test_elec_energy_noalloc_matches_existing_entry_point():
    from pyseqm_b200 import _lib
    L = _lib.lib()
    dev = torch.device('cuda:0')
    torch.cuda.set_device(dev)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    # call 1: many small molecules (one chunk each); call 2, right after, a few large ones (several chunks each) with
    # the XL-BOMD form of the energy and an `active` mask: the second call finds the first call's partial sums in the
    # shared static array and must not be affected by them
    for nmol, molsize, seed, xl, masked in ((300, 12, 1, 0, False), (5, 70, 2, 1, True), (300, 12, 3, 0, True)):
        P, F, D, M = _inputs(nmol, molsize, seed, dev)
        active = None
        if masked:
            active = torch.ones(nmol, dtype=torch.uint8, device=dev)
            active[1::2] = 0
        e_ref, e_new = _both(L, nmol, molsize, P, F, D, M, xl, active, stream)
        scale = max(1.0, float(np.abs(e_ref).max()))
        assert np.abs(e_new - e_ref).max() <= TOL * scale, (nmol, molsize, np.abs(e_new - e_ref).max())
        if masked:      # inactive molecules are left untouched by both entry points
            assert (e_new[1::2] == -7.0).all() and (e_ref[1::2] == -7.0).all()
        assert np.isfinite(e_new).all() and np.abs(e_new[0::2] + 7.0).min() > 0.0      # really written


def test_elec_energy_noalloc_rejects_sizes_beyond_the_static_scratch():
    from pyseqm_b200 import _lib
    L = _lib.lib()
    dev = torch.device('cuda:0')
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    dummy = torch.zeros(16, dtype=torch.float64, device=dev)
    # 70,000 molecules x 1 chunk > 65,536 partial sums: refused before anything is launched or dereferenced
    rc = L.seqm_elec_energy_noalloc(70000, 4, _p(dummy), _p(dummy), _p(dummy), _p(None), 0, _p(None), _p(dummy), stream)
    assert rc == -1
    msg = L.seqm_last_error().decode()
    assert 'SEQM_NOALLOC_MAX_PARTIALS' in msg and '70000' in msg
    torch.cuda.synchronize()
