"""Builds tests/hostcheck/hostcheck.cpp (the kernels' templated pair mathematics compiled for the host)
and wraps it with ctypes.  Test infrastructure only."""
import ctypes
import os
import subprocess
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_lib = None


def lib():
    global _lib
    if _lib is None:
        out = os.path.join(tempfile.gettempdir(), 'libseqm_hostcheck_%d.so' % os.getuid())
        src = os.path.join(HERE, 'hostcheck', 'hostcheck.cpp')
        subprocess.check_call(['g++', '-O1', '-std=c++17', '-shared', '-fPIC', '-x', 'c++', src, '-o', out])
        _lib = ctypes.CDLL(out)
        _lib.hc_atompar_doubles.restype = ctypes.c_int
    return _lib


def atompar(zs, zp, bs, bp, dd, qq, rho0, rho1, rho2, alpha, tore, K=(), L=(), M=()):
    v = np.zeros(lib().hc_atompar_doubles())
    v[0:4] = [zs, zp, bs, bp]
    v[4:9] = [dd, qq, rho0, rho1, rho2]
    v[9:11] = [alpha, tore]
    v[11:14] = [zs ** 1.5, zs ** 2.5, zp ** 2.5]
    for k in range(len(K)):
        v[14 + k] = K[k]
        v[18 + k] = L[k]
        v[22 + k] = M[k]
    return v


def _p(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_double))


def pair(ni, nj, x, r, apar, bpar, ngauss):
    w = np.zeros(100)
    di = np.zeros(16)
    e = np.zeros(1)
    x = np.ascontiguousarray(x, dtype=np.float64)
    lib().hc_pair(int(ni), int(nj), _p(x), ctypes.c_double(r), _p(apar), _p(bpar), int(ngauss), _p(w), _p(di), _p(e))
    return w.reshape(10, 10), di.reshape(4, 4), e[0]


def pair_dual(ni, nj, x, r, apar, bpar, ngauss):
    w = np.zeros(500)
    di = np.zeros(80)
    e = np.zeros(5)
    x = np.ascontiguousarray(x, dtype=np.float64)
    lib().hc_pair_dual(int(ni), int(nj), _p(x), ctypes.c_double(r), _p(apar), _p(bpar), int(ngauss), _p(w), _p(di), _p(e))
    return w.reshape(10, 10, 5), di.reshape(4, 4, 5), e


def rotate_xx_adjoint(x, r, apar, bpar, G):
    out = np.zeros(4)
    x = np.ascontiguousarray(x, dtype=np.float64)
    G = np.ascontiguousarray(G, dtype=np.float64)
    lib().hc_rotate_xx_adjoint(_p(x), ctypes.c_double(r), _p(apar), _p(bpar), _p(G), _p(out))
    return out


def overlap_xx_adjoint(x, r, apar, bpar, F):
    out = np.zeros(4)
    x = np.ascontiguousarray(x, dtype=np.float64)
    F = np.ascontiguousarray(F, dtype=np.float64)
    lib().hc_overlap_xx_adjoint(_p(x), ctypes.c_double(r), _p(apar), _p(bpar), _p(F), _p(out))
    return out
