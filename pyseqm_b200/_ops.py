"""Thin op layer over the C ABI: allocates outputs with torch, passes raw device pointers and the
current CUDA stream, and wraps the pair pass + its hand-written adjoint as a torch.autograd.Function so
`Force`'s `L.backward()` and `ForceXL`'s `grad(L, coordinates)` work as in the reference."""
import ctypes

import torch

from . import _lib
from ._lib import ScfOptions, WStore, check, lib
from ._topology import PairTopology

N_GAUSS = {'MNDO': 0, 'AM1': 4, 'PM3': 2}
LAUNCHES = [0]      # C-ABI calls issued (bench.py reports kernel launches from its own counters)


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _p(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


def _f64(t):
    if t.dtype != torch.float64:
        raise TypeError("pyseqm_b200 computes in float64 only (got %s); call torch.set_default_dtype(torch.float64)" % t.dtype)
    if not t.is_cuda:
        raise RuntimeError("pyseqm_b200 runs on CUDA tensors only (there is no CPU path)")
    return t.detach().contiguous()


class WTensors:
    """class-tiled two-centre integrals"""

    def __init__(self, topo):
        n0, n1, n2 = topo.w_sizes()
        dev = topo.device
        self.hh = torch.empty(n0, dtype=torch.float64, device=dev)
        self.xh = torch.empty(n1, dtype=torch.float64, device=dev)
        self.xx = torch.empty(n2, dtype=torch.float64, device=dev)
        self.c = WStore(self.hh.data_ptr(), self.xh.data_ptr(), self.xx.data_ptr())


def atom_pack(topo, par, method):
    """parameters dict (per-atom vectors, reference basics.py:148-154) -> packed per-atom records"""
    ng = N_GAUSS[method]
    A = topo.natoms
    out = torch.empty(A * lib().seqm_atompar_bytes(), dtype=torch.uint8, device=topo.device)
    g = [None, None, None]
    if ng:
        for q, c in enumerate('KLM'):
            g[q] = torch.stack([_f64(par['Gaussian%d_%s' % (k + 1, c)]) for k in range(ng)], dim=1).contiguous()
    v = [_f64(par[k]) for k in ('zeta_s', 'zeta_p', 'beta_s', 'beta_p', 'g_ss', 'g_pp', 'g_p2', 'h_sp', 'alpha')]
    check(lib().seqm_atom_pack(A, _p(topo.Z), *[_p(x) for x in v], _p(g[0]), _p(g[1]), _p(g[2]), ng, _p(out),
                               _stream()), 'seqm_atom_pack')
    return out


def hcore(topo, atompar, xij, rij, U_ss, U_pp, method, want_enuc=True):
    """-> (WTensors, M[B*S*S,4,4], EnucAB[P] or None, pairvec scratch)"""
    dev = topo.device
    w = WTensors(topo)
    M = torch.zeros(topo.nmol * topo.molsize * topo.molsize, 4, 4, dtype=torch.float64, device=dev)
    en = torch.empty(topo.npairs, dtype=torch.float64, device=dev) if want_enuc else None
    pv = torch.empty(max(topo.npairs, 1) * 20, dtype=torch.float64, device=dev)
    xij, rij, U_ss, U_pp = _f64(xij), _f64(rij), _f64(U_ss), _f64(U_pp)
    check(lib().seqm_hcore(topo.ref(), _p(atompar), _p(xij), _p(rij), _p(U_ss), _p(U_pp), N_GAUSS[method], w.c,
                           _p(M), _p(en), _p(pv), _stream()), 'seqm_hcore')
    return w, M, en, pv


def w_expand(topo, w):
    out = torch.empty(topo.npairs, 10, 10, dtype=torch.float64, device=topo.device)
    check(lib().seqm_w_expand(topo.ref(), w.c, _p(out), _stream()), 'seqm_w_expand')
    return out


def fock(topo, w, P, M, par5, pv=None, F=None, active=None, want_energy=False):
    """par5 = (g_ss, g_pp, g_sp, g_p2, h_sp) per-atom vectors.  -> F (and Eelec)"""
    n4 = 4 * topo.molsize
    if F is None:
        F = torch.zeros(topo.nmol, n4, n4, dtype=torch.float64, device=topo.device)
    if pv is None:
        pv = torch.empty(max(topo.npairs, 1) * 20, dtype=torch.float64, device=topo.device)
    E = torch.zeros(topo.nmol, dtype=torch.float64, device=topo.device) if want_energy else None
    P, M = _f64(P), _f64(M)
    g = [_f64(x) for x in par5]
    work = _energy_work(topo.nmol, topo.molsize, topo.device) if want_energy else None
    check(lib().seqm_fock(topo.ref(), w.c, _p(P), _p(M), *[_p(x) for x in g], _p(active), _p(F), _p(E), _p(pv),
                          _p(work), _stream()), 'seqm_fock')
    return (F, E) if want_energy else F


def _energy_work(nmol, molsize, device):
    """scratch of the chunked energy sum (large molecules only; None keeps the one-CTA-per-molecule kernel)"""
    if 4 * molsize <= 256:
        return None
    return torch.empty(lib().seqm_elec_energy_workspace(nmol, molsize), dtype=torch.uint8, device=device)


def elec_energy(P, F, M, nmol, molsize, D=None):
    E = torch.zeros(nmol, dtype=torch.float64, device=P.device)
    P, F, M = _f64(P), _f64(F), _f64(M)
    Dc = _f64(D) if D is not None else None
    work = _energy_work(nmol, molsize, P.device)
    check(lib().seqm_elec_energy(nmol, molsize, _p(P), _p(F), _p(M), _p(Dc), 1 if D is not None else 0, _p(None),
                                 _p(E), _p(work), _stream()), 'seqm_elec_energy')
    return E


def eigh_density(F, nheavy, nhydro, nocc, molsize, mmax, want_eig=False, active=None, info=None):
    """-> P (and e[B,n], v[B,n,n] with the packed eigenvectors in the top-left m x m corner)"""
    F = _f64(F)
    B, n4 = F.shape[0], F.shape[1]
    P = torch.zeros_like(F)
    e = torch.zeros(B, n4, dtype=torch.float64, device=F.device) if want_eig else None
    v = torch.zeros_like(F) if want_eig else None
    check(lib().seqm_eigh_density(B, molsize, int(mmax), _p(F), _p(nheavy), _p(nhydro), _p(nocc), _p(active), _p(P),
                                  _p(e), _p(v), _p(info), _stream()), 'seqm_eigh_density')
    return (P, e, v) if want_eig else P


def sp2_density(F, nheavy, nhydro, nocc, molsize, mmax, eps, active=None, want_niter=False):
    if _sp2_shard['ranks'] >= 1 or _sp2_shard['comm'] is not None:
        return sp2_density_sharded(F, nheavy, nhydro, nocc, molsize, mmax, eps, _sp2_shard['comm'], _sp2_shard['ranks'],
                                   active, want_niter)
    F = _f64(F)
    B = F.shape[0]
    D = torch.zeros_like(F)
    work = torch.empty(lib().seqm_sp2_workspace(B, molsize), dtype=torch.uint8, device=F.device)
    nit = torch.zeros(B, dtype=torch.int32, device=F.device) if want_niter else None
    check(lib().seqm_sp2_density(B, molsize, int(mmax), _p(F), _p(nheavy), _p(nhydro), _p(nocc), float(eps),
                                 _p(active), _p(D), _p(nit), _p(work), _stream()), 'seqm_sp2_density')
    return (D, nit) if want_niter else D


def sp2_density_async(F, nheavy, nhydro, nocc, molsize, mmax, eps, steps, work=None, D=None, unfinished=None):
    """fixed number of purification steps, no host synchronisation (CUDA-graph capturable); unfinished (int32[1] on the
    device) receives the number of molecules that would have needed more steps"""
    F = _f64(F)
    B = F.shape[0]
    if D is None:
        D = torch.zeros_like(F)
    if work is None:
        work = torch.empty(lib().seqm_sp2_workspace(B, molsize), dtype=torch.uint8, device=F.device)
    if unfinished is None:
        unfinished = torch.zeros(1, dtype=torch.int32, device=F.device)
    check(lib().seqm_sp2_density_async(B, molsize, int(mmax), _p(F), _p(nheavy), _p(nhydro), _p(nocc), float(eps), _p(None),
                                       _p(D), _p(None), _p(work), int(steps), _p(unfinished), _stream()),
          'seqm_sp2_density_async')
    return D, unfinished


# ---- tile-sharded SP2 (single large system over the GPUs of one box; include/seqm_b200.h, sp2_kernels.cu) ----
_sp2_shard = {'comm': None, 'ranks': 0}      # handle of seqm_nccl_comm_create / number of (emulated) ranks


def sp2_sharding(comm=None, ranks=0):
    """route sp2_density() and the SP2 inside scf() through seqm_sp2_density_nccl: comm = handle from
    nccl_comm_create(), or comm None with ranks >= 1 for the single-GPU emulation; (None, 0) switches it off"""
    _sp2_shard['comm'], _sp2_shard['ranks'] = comm, int(ranks)
    lib().seqm_scf_sp2_comm(comm, int(ranks))


def nccl_unique_id():
    buf = ctypes.create_string_buffer(128)
    check(lib().seqm_nccl_unique_id(buf), 'seqm_nccl_unique_id')
    return buf.raw


def nccl_comm_create(uid, rank, nranks):
    h = ctypes.c_void_p()
    check(lib().seqm_nccl_comm_create(ctypes.create_string_buffer(bytes(uid), 128), int(rank), int(nranks),
                                      ctypes.byref(h)), 'seqm_nccl_comm_create')
    return h


def nccl_comm_destroy(h):
    if _sp2_shard['comm'] is h:
        sp2_sharding(None, 0)
    check(lib().seqm_nccl_comm_destroy(h), 'seqm_nccl_comm_destroy')


def sp2_density_sharded(F, nheavy, nhydro, nocc, molsize, mmax, eps, comm=None, ranks=1, active=None, want_niter=False):
    F = _f64(F)
    B = F.shape[0]
    D = torch.zeros_like(F)
    nr = int(ranks) if comm is None else 64      # workspace bound; the library knows the communicator's size
    work = torch.empty(lib().seqm_sp2_nccl_workspace(B, molsize, nr), dtype=torch.uint8, device=F.device)
    nit = torch.zeros(B, dtype=torch.int32, device=F.device) if want_niter else None
    check(lib().seqm_sp2_density_nccl(comm, int(ranks), B, molsize, int(mmax), _p(F), _p(nheavy), _p(nhydro), _p(nocc),
                                      float(eps), _p(active), _p(D), _p(nit), _p(work), _stream()),
          'seqm_sp2_density_nccl')
    return (D, nit) if want_niter else D


def scf_is_fused(topo, sp2, debug):
    """the one-launch form of seqm_scf applies (include/seqm_b200.h)"""
    return (not sp2[0]) and (not debug) and ((max(int(topo.mmax), 1) + 1) & ~1) <= 32


def scf(topo, w, M, par5, P0, eps, converger, sp2, max_iter, debug, pv, status=None):
    """-> P, F, Eelec, notconverged(bool), niter[B], (iterations, left).
    status: pinned int64[2] tensor -> no synchronisation (one-launch form only); the last element is then None and
    status receives {left, iterations} when the stream gets there (see ScfStatus)."""
    dev = topo.device
    B, S = topo.nmol, topo.molsize
    n4 = 4 * S
    opt = ScfOptions()
    opt.converger = int(converger[0])
    opt.alpha = float(converger[1]) if (len(converger) > 1 and int(converger[0]) == 0) else 0.0
    opt.use_sp2 = 1 if sp2[0] else 0
    opt.sp2_eps = float(sp2[1]) if sp2[0] else 0.0
    opt.eps = float(eps)
    opt.max_iter = int(max_iter)
    opt.debug = 1 if debug else 0
    if opt.converger not in (0, 1, 2):
        raise ValueError("scf_converger should be [0, alpha], [1] or [2]")
    from .seqm_functions.diag import apply_check_degeneracy
    apply_check_degeneracy(topo.mmax, sp2=bool(sp2[0]))
    use_async = status is not None and scf_is_fused(topo, sp2, debug)
    opt.async_status = 1 if use_async else 0
    P = _f64(P0).clone()
    F = torch.empty(B, n4, n4, dtype=torch.float64, device=dev)
    E = torch.zeros(B, dtype=torch.float64, device=dev)
    nc = torch.ones(B, dtype=torch.uint8, device=dev)
    nit = torch.zeros(B, dtype=torch.int32, device=dev)
    work = torch.empty(lib().seqm_scf_workspace(B, S, ctypes.byref(opt)), dtype=torch.uint8, device=dev)
    iters = (ctypes.c_int32 * 2)(0, 0)
    out_ptr = ctypes.c_void_p(status.data_ptr()) if use_async else ctypes.cast(iters, ctypes.c_void_p)
    g = [_f64(x) for x in par5]
    check(lib().seqm_scf(topo.ref(), w.c, _p(_f64(M)), *[_p(x) for x in g], _p(topo.nheavy), _p(topo.nhydro),
                         _p(topo.nocc), int(topo.mmax), ctypes.byref(opt), _p(P), _p(F), _p(E), _p(nc), _p(nit),
                         _p(pv), _p(work), out_ptr, _stream()), 'seqm_scf')
    if use_async:
        return P, F, E, nc.bool(), nit, None
    return P, F, E, nc.bool(), nit, (int(iters[0]), int(iters[1]))


def pair_grad(topo, atompar, xij, rij, method, P, D, gE, gNuc):
    gx = torch.empty(topo.npairs, 3, dtype=torch.float64, device=topo.device)
    gr = torch.empty(topo.npairs, dtype=torch.float64, device=topo.device)
    xij, rij, P = _f64(xij), _f64(rij), _f64(P)
    Dc = _f64(D) if D is not None else None
    gE = _f64(gE) if gE is not None else None
    gNuc = _f64(gNuc) if gNuc is not None else None
    check(lib().seqm_pair_grad(topo.ref(), _p(atompar), _p(xij), _p(rij), N_GAUSS[method], _p(P), _p(Dc), _p(gE),
                               _p(gNuc), _p(gx), _p(gr), _stream()), 'seqm_pair_grad')
    return gx, gr


class PairGeometry(torch.autograd.Function):
    """coordinates[B,S,3] -> (xij[P,3], rij[P]): the Parser's geometry (reference basics.py:91-108) as one
    kernel, with the per-atom fixed-order adjoint kernel as backward."""

    @staticmethod
    def forward(ctx, coordinates, topo, lcf):
        R = _f64(coordinates)
        xij = torch.empty(topo.npairs, 3, dtype=torch.float64, device=R.device)
        rij = torch.empty(topo.npairs, dtype=torch.float64, device=R.device)
        check(lib().seqm_pair_geometry(topo.ref(), _p(R), float(lcf), _p(xij), _p(rij), _stream()), 'seqm_pair_geometry')
        ctx.topo, ctx.lcf, ctx.shape = topo, float(lcf), coordinates.shape
        ctx.save_for_backward(xij, rij)
        return xij, rij

    @staticmethod
    def backward(ctx, gx, gr):
        xij, rij = ctx.saved_tensors
        gR = torch.zeros(ctx.shape, dtype=torch.float64, device=xij.device)
        gx, gr = _f64(gx), _f64(gr)
        check(lib().seqm_pair_geometry_grad(ctx.topo.ref(), _p(xij), _p(rij), _p(gx), _p(gr), ctx.lcf, _p(gR), _stream()),
              'seqm_pair_geometry_grad')
        return gR, None, None


class PairEnergy(torch.autograd.Function):
    """(xij, rij) -> (Eelec[B], EnucAB[P]) at fixed density.  The values were computed by the forward
    kernels; backward is the fused adjoint kernel (seqm_pair_grad), i.e. exactly the part of the graph the
    reference's autograd replays with SCF0 treating P as a constant (scf_loop.py:661-668)."""

    @staticmethod
    def forward(ctx, xij, rij, state):
        ctx.state = state
        ctx.save_for_backward(xij, rij)
        return state['Eelec'].clone(), state['EnucAB'].clone()

    @staticmethod
    def backward(ctx, gE, gNuc):
        xij, rij = ctx.saved_tensors
        s = ctx.state
        gx, gr = pair_grad(s['topo'], s['atompar'], xij, rij, s['method'], s['P'], s.get('D'), gE, gNuc)
        return gx, gr, None


def dgemm_batched(A, B):
    A, B = _f64(A), _f64(B)
    nb, n = A.shape[0], A.shape[1]
    C = torch.zeros_like(A)
    check(lib().seqm_dgemm_batched(n, n, _p(A), _p(B), _p(C), n * n, nb, _p(None), _stream()), 'seqm_dgemm_batched')
    return C
