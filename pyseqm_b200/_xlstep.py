"""One XL-BOMD step (reference XLBOMD.py:277-307) as a fixed sequence of libseqm_b200 launches on preallocated
buffers: Verlet kick + drift, density extrapolation, pair geometry, pair pass, Fock build, ONE density solve (SP2 with
a fixed number of purification steps, or eigh), shadow energy, adjoint pass, geometry adjoint, energies, Verlet kick.
No ATen kernel, no allocation and no host synchronisation inside a step, so the step can be captured in ONE CUDA graph
(`graph=True`) and replayed; the MD step index lives on the device.

The state tensors (coordinates, velocities, acc, D, P, Pt) are updated IN PLACE, like `coordinates` and `velocities`
already are in the reference's one_step; the energies / forces returned by step() are views of buffers that the next
step overwrites.  Applies when the species-only data can be cached (`_engine.Prepared`: all-pairs cutoff, no per-call
parameters); XL_BOMD.one_step falls back to the autograd form otherwise."""
import ctypes
import warnings

import torch

from . import _ops
from ._lib import WStore, check, lib

_p = _ops._p


class XLStep:
    def __init__(self, md, const, prep, species, coordinates, velocities, acc, D, P, Pt, mass, step0, graph=False,
                 sp2_steps=None):
        dev = coordinates.device
        self.md, self.prep, self.topo = md, prep, prep.topo
        self.method = md.conservative_force.energy.method
        self.sp2 = list(md.seqm_parameters['sp2'])
        topo = self.topo
        B, S = topo.nmol, topo.molsize
        n4 = 4 * S
        f64 = dict(dtype=torch.float64, device=dev)
        self.B, self.S = B, S
        # state (shared with the caller, updated in place)
        self.x, self.v, self.acc, self.D, self.P, self.Pt = coordinates.detach(), velocities, acc, D, P, Pt
        for t in (self.x, self.v, self.acc, self.D, self.P, self.Pt):
            if not (t.is_cuda and t.dtype == torch.float64 and t.is_contiguous()):
                raise ValueError("XLStep needs contiguous CUDA float64 state tensors")
        self.mass = mass.reshape(B, S).contiguous()
        self.step_dev = torch.full((1,), int(step0), dtype=torch.int64, device=dev)
        self.coeff = (ctypes.c_double * (2 * md.m))(*[float(c) for c in md.coeff.detach().cpu().tolist()])
        # buffers
        self.xij = torch.empty(topo.npairs, 3, **f64)
        self.rij = torch.empty(topo.npairs, **f64)
        self._side = torch.cuda.Stream(device=dev)      # second branch of the step (density extrapolation)
        self.w = _ops.WTensors(topo)
        self.M = torch.zeros(B * S * S, 4, 4, **f64)          # only the pair pattern is ever written: zero once
        self.EnucAB = torch.empty(topo.npairs, **f64)
        self.pv = torch.empty(max(topo.npairs, 1) * 20, **f64)
        self.F = torch.zeros(B, n4, n4, **f64)
        self.Eelec = torch.zeros(B, **f64)
        self.ework = _ops._energy_work(B, S, dev)
        self.gx = torch.empty(topo.npairs, 3, **f64)
        self.gr = torch.empty(topo.npairs, **f64)
        self.gR = torch.zeros(B, S, 3, **f64)
        self.force = torch.zeros(B, S, 3, **f64)
        self.Enuc, self.Etot, self.Hf = (torch.zeros(B, **f64) for _ in range(3))
        par = prep.parameters
        self.par5 = [_ops._f64(par[k]) for k in ('g_ss', 'g_pp', 'g_sp', 'g_p2', 'h_sp')]
        self.U_ss, self.U_pp = _ops._f64(par['U_ss']), _ops._f64(par['U_pp'])
        # per-molecule pair ranges and the constant part of the heat of formation (energy.py:109-118)
        ar = torch.arange(B + 1, device=dev, dtype=prep.pair_molid.dtype)
        self.pair_off = torch.searchsorted(prep.pair_molid.contiguous(), ar).to(torch.int32).contiguous()
        zero = torch.zeros(B, **f64)
        hc = -zero.index_add(0, prep.atom_molid, prep.Eiso)
        if md.conservative_force.energy.Hf_flag:
            hc = hc + zero.index_add(0, prep.atom_molid, const.eheat[prep.Z].to(torch.float64))
        self.hconst = hc.contiguous()
        # density solver
        self.use_sp2 = bool(self.sp2[0])
        if self.use_sp2:
            self.sp2_work = torch.empty(lib().seqm_sp2_workspace(B, S), dtype=torch.uint8, device=dev)
            self.unfinished = torch.zeros(1, dtype=torch.int32, device=dev)
            self.unfinished_host = torch.zeros(1, dtype=torch.int32).pin_memory()
            self.sp2_steps = int(sp2_steps) if sp2_steps else 0
        self.graph = None
        self.want_graph = bool(graph)
        self.nsteps = 0
        self.launches_per_step = 0
        self._event = None

    # ---- the launches of one step, on the current stream
    def _enqueue(self):
        L, topo, md = lib(), self.topo, self.md
        st = _ops._stream()
        n3 = self.x.numel()
        dt = float(md.timestep)
        # fork: the density extrapolation (reads D and the ring Pt of the previous steps, HBM-bound) does not depend on the
        # new coordinates, so it runs on a side stream next to the Verlet drift, the pair geometry and the pair pass
        # (FP64-latency-bound); joined before the Fock build.  Inside a captured graph these are two parallel branches.
        main = torch.cuda.current_stream()
        side = self._side
        side.wait_stream(main)
        with torch.cuda.stream(side):
            check(L.seqm_xl_extrapolate(self.D.numel(), md.m, float(md.coeff_D), self.coeff, _p(self.step_dev), 0,
                                        _p(self.D), _p(self.Pt), _p(self.P), _ops._stream()), 'seqm_xl_extrapolate')
        check(L.seqm_verlet_kick_drift(n3, dt, _p(self.acc), _p(self.v), _p(self.x), st), 'seqm_verlet_kick_drift')
        check(L.seqm_pair_geometry(topo.ref(), _p(self.x), self.prep.lcf, _p(self.xij), _p(self.rij), st), 'seqm_pair_geometry')
        ng = _ops.N_GAUSS[self.method]
        check(L.seqm_hcore(topo.ref(), _p(self.prep.atompar), _p(self.xij), _p(self.rij), _p(self.U_ss), _p(self.U_pp), ng,
                           self.w.c, _p(self.M), _p(self.EnucAB), _p(self.pv), st), 'seqm_hcore')
        main.wait_stream(side)
        check(L.seqm_fock(topo.ref(), self.w.c, _p(self.P), _p(self.M), *[_p(t) for t in self.par5], _p(None), _p(self.F),
                          _p(None), _p(self.pv), _p(None), st), 'seqm_fock')
        if self.use_sp2:
            check(L.seqm_sp2_density_async(self.B, self.S, int(topo.mmax), _p(self.F), _p(topo.nheavy), _p(topo.nhydro),
                                           _p(topo.nocc), float(self.sp2[1]), _p(None), _p(self.D), _p(None),
                                           _p(self.sp2_work), self.sp2_steps, _p(self.unfinished), st),
                  'seqm_sp2_density_async')
            self.unfinished_host.copy_(self.unfinished, non_blocking=True)
        else:
            check(L.seqm_eigh_density(self.B, self.S, int(topo.mmax), _p(self.F), _p(topo.nheavy), _p(topo.nhydro),
                                      _p(topo.nocc), _p(None), _p(self.D), _p(None), _p(None), _p(None), st),
                  'seqm_eigh_density')
        check(L.seqm_elec_energy(self.B, self.S, _p(self.P), _p(self.F), _p(self.M), _p(self.D), 1, _p(None),
                                 _p(self.Eelec), _p(self.ework), st), 'seqm_elec_energy')
        check(L.seqm_pair_grad(topo.ref(), _p(self.prep.atompar), _p(self.xij), _p(self.rij), ng, _p(self.P), _p(self.D),
                               _p(None), _p(None), _p(self.gx), _p(self.gr), st), 'seqm_pair_grad')
        check(L.seqm_pair_geometry_grad(topo.ref(), _p(self.xij), _p(self.rij), _p(self.gx), _p(self.gr), self.prep.lcf,
                                        _p(self.gR), st), 'seqm_pair_geometry_grad')
        check(L.seqm_energy_finish(self.B, _p(self.pair_off), _p(self.EnucAB), _p(self.Eelec), _p(self.hconst),
                                   _p(self.Enuc), _p(self.Etot), _p(self.Hf), st), 'seqm_energy_finish')
        check(L.seqm_verlet_finish(n3, dt, float(md.acc_scale), _p(self.gR), _p(self.mass), _p(self.force), _p(self.acc),
                                   _p(self.v), _p(self.step_dev), st), 'seqm_verlet_finish')

    def _calibrate_sp2(self):
        """purification steps the fixed-count form must cover: what the synchronising form needs on the current state
        (one throw-away solve on a copy of F), plus a margin for the drift of the spectrum along the trajectory"""
        topo = self.topo
        _, nit = _ops.sp2_density(self.F.clone(), topo.nheavy, topo.nhydro, topo.nocc, self.S, topo.mmax, self.sp2[1],
                                  want_niter=True)
        need = int(nit.max())
        self.sp2_steps = min(64, ((need + 4 + 1) // 2) * 2)

    def _check_unfinished(self):
        """every SP2 solve must finish inside the fixed number of purification steps: the device keeps a running total
        of the molecules that did not, mirrored into pinned memory by every step; it is looked at without blocking"""
        if not self.use_sp2 or self._event is None or not self._event.query():
            return
        left = int(self.unfinished_host[0])
        if left:
            warnings.warn("XL-BOMD: SP2 needed more than the %d purification steps of the fixed-count step for %d "
                          "molecule-solve(s) up to step %d; the count is raised for the following steps"
                          % (self.sp2_steps, left, self.nsteps))
            self.sp2_steps = min(64, self.sp2_steps + 4)
            self.graph = None
            torch.cuda.current_stream().synchronize()
            self.unfinished.zero_()
            self.unfinished_host.zero_()

    def step(self):
        """advance the state by one MD step; -> (Hf, force) views"""
        if self.use_sp2 and not self.sp2_steps:
            # first step: F is not known yet -> run this step's launches once with a generous count, then calibrate
            self.sp2_steps = 64
            self._enqueue()
            self._calibrate_sp2()
        elif self.want_graph:
            self._check_unfinished()
            if self.graph is None:
                torch.cuda.synchronize()
                g = torch.cuda.CUDAGraph()
                l0 = lib().seqm_launch_count()
                with torch.cuda.graph(g):          # capturing does not execute: the state advances only in replay()
                    self._enqueue()
                self.launches_per_step = int(lib().seqm_launch_count() - l0)      # kernels one replay launches
                self.graph = g
            self.graph.replay()
        else:
            self._check_unfinished()
            self._enqueue()
        if self.use_sp2:
            self._event = torch.cuda.Event()
            self._event.record()
        self.nsteps += 1
        return self.Hf, self.force

    def matches(self, coordinates, velocities, acc, D, P, Pt):
        same = lambda a, b: a.data_ptr() == b.data_ptr() and a.shape == b.shape
        return (same(coordinates, self.x) and same(velocities, self.v) and same(acc, self.acc) and same(D, self.D)
                and same(P, self.P) and same(Pt, self.Pt))
