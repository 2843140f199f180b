"""Parser, Pack_Parameters, Hamiltonian, Energy, Force — the API of the reference's seqm/basics.py (same
constructor dicts, call signatures and return tuples), with the numerical work done by the sm_100a kernels
of libseqm_b200.so.  Tensors must be CUDA float64."""
import os
import time

import torch

from . import _engine, _ops
from .seqm_functions import scf_loop as _scf_loop_mod
from .seqm_functions.constants import ev
from .seqm_functions.energy import elec_energy_isolated_atom, heat_formation, total_energy
from .seqm_functions.parameters import params
from .seqm_functions.scf_loop import scf_loop

_BASE = ['U_ss', 'U_pp', 'zeta_s', 'zeta_p', 'beta_s', 'beta_p', 'g_ss', 'g_sp', 'g_pp', 'g_p2', 'h_sp', 'alpha']


def _gauss(n):
    return ['Gaussian%d_%s' % (k, c) for c in 'KLM' for k in range(1, n + 1)]


parameterlist = {'AM1': _BASE + _gauss(4), 'MNDO': list(_BASE), 'PM3': _BASE + _gauss(2)}


class Parser(torch.nn.Module):
    """species/coordinates -> the 16 index and geometry tensors of reference basics.py:43-118.
    Pairs are ordered (molecule, a, b) with a < b, exactly like the reference's boolean-mask selection;
    the arithmetic on coordinates uses the same operations so xij and rij agree bit for bit."""

    def __init__(self, seqm_parameters):
        super().__init__()
        self.outercutoff = seqm_parameters['pair_outer_cutoff']
        self.elements = seqm_parameters['elements']
        # pyseqm_b200 extension: 'parser': 'torch' keeps the index arithmetic below on CUDA tensors too (comparison /
        # test hook); the default on CUDA float64 inputs is the device kernel pair of _parser.device_parse
        self.backend = seqm_parameters.get('parser', 'device')
        self.last_topology = None

    def forward(self, constansts, species, coordinates, *args, **kwargs):
        self.last_topology = None
        if (self.backend == 'device' and coordinates.is_cuda and coordinates.dtype == torch.float64
                and species.dtype == torch.int64 and not args):
            from ._parser import device_parse
            out, topo = device_parse(constansts, species, coordinates, self.outercutoff, kwargs.get('charges'))
            self.last_topology = (out[12], topo)          # keyed by the idxi tensor it belongs to
            return out
        device = coordinates.device
        nmol, molsize = species.shape
        nonblank = species > 0
        flat = nonblank.reshape(-1)
        real_atoms = torch.nonzero(flat).reshape(-1)
        n_real = real_atoms.shape[0]
        inv_real_atoms = torch.zeros((nmol * molsize,), device=device, dtype=torch.int64)
        inv_real_atoms[real_atoms] = torch.arange(n_real, device=device, dtype=torch.int64)
        Z = species.reshape(-1)[real_atoms]
        nHeavy = torch.sum(species > 1, dim=1)
        nHydro = torch.sum(species == 1, dim=1)
        n_charge = torch.sum(constansts.tore[species], dim=1).reshape(-1).type(torch.int64)
        if 'charges' in kwargs and torch.is_tensor(kwargs['charges']):
            n_charge = n_charge - kwargs['charges'].reshape(-1).type(torch.int64)
        nocc = torch.div(n_charge, 2, rounding_mode='floor')
        if ((n_charge % 2) == 1).any():
            raise ValueError("Only closed shell system (with even number of electrons) are supported")
        mol_of = torch.arange(nmol, device=device, dtype=torch.int64).repeat_interleave(molsize)
        slot_of = torch.arange(molsize, device=device, dtype=torch.int64).repeat(nmol)
        maskd = (mol_of * molsize * molsize + slot_of * (molsize + 1))[real_atoms]
        atom_molid = mol_of[real_atoms]
        # all a < b slot pairs, row-major = (a, b) lexicographic
        a, b = torch.triu_indices(molsize, molsize, offset=1, device=device)
        npm = a.shape[0]
        paircoord_raw = coordinates[:, b, :] - coordinates[:, a, :]            # R_b - R_a
        pairdist2 = (paircoord_raw ** 2).sum(dim=2)
        pairdist2 = torch.where(pairdist2 == 0.0, torch.full_like(pairdist2, 1.0e-4), pairdist2)
        pairdist_raw = torch.sqrt(pairdist2)
        keep = (nonblank[:, a] & nonblank[:, b] & (pairdist_raw < self.outercutoff)).reshape(-1)
        paircoord = paircoord_raw.reshape(-1, 3)[keep]
        pairdist = pairdist_raw.reshape(-1)[keep]
        rij = pairdist * constansts.length_conversion_factor
        molp = torch.arange(nmol, device=device, dtype=torch.int64).unsqueeze(1).expand(nmol, npm).reshape(-1)[keep]
        first = molp * molsize + a.repeat(nmol)[keep]
        second = molp * molsize + b.repeat(nmol)[keep]
        idxi = inv_real_atoms[first]
        idxj = inv_real_atoms[second]
        ni = Z[idxi]
        nj = Z[idxj]
        xij = paircoord / pairdist.unsqueeze(1)
        mask = first * molsize + second % molsize
        pair_molid = molp
        return nmol, molsize, nHeavy, nHydro, nocc, Z, maskd, atom_molid, mask, pair_molid, ni, nj, idxi, idxj, xij, rij


class Pack_Parameters(torch.nn.Module):
    """per-atom parameter vectors p[Z, i] (reference basics.py:120-154)"""

    def __init__(self, seqm_parameters):
        super().__init__()
        self.elements = seqm_parameters['elements']
        self.learned_list = seqm_parameters['learned']
        self.method = seqm_parameters['method']
        self.filedir = seqm_parameters['parameter_file_dir'] if 'parameter_file_dir' in seqm_parameters \
            else os.path.abspath(os.path.dirname(__file__)) + '/params/'
        self.parameters = parameterlist[self.method]
        self.required_list = [name for name in self.parameters if name not in self.learned_list]
        self.nrp = len(self.required_list)
        self.p = params(method=self.method, elements=self.elements, root_dir=self.filedir,
                        parameters=self.required_list)

    def forward(self, Z, learned_params=dict()):
        # a NEW dict: the reference fills the caller's dict in place (basics.py:148-154), which with the shared default
        # argument `learned_parameters=dict()` of Energy / Force leaves the module-level default non-empty after the first
        # call - here that would switch the cached-species path off for good and let BucketedBatch's worker threads
        # overwrite each other's parameters
        out = dict(learned_params)
        for i in range(self.nrp):
            out[self.required_list[i]] = self.p[Z, i]
        return out


class Hamiltonian(torch.nn.Module):
    """SCF (reference basics.py:156-251): -> F, e, P, Hcore, w, charge, notconverged"""

    def __init__(self, seqm_parameters):
        super().__init__()
        self.eps = torch.nn.Parameter(torch.as_tensor(seqm_parameters['scf_eps']), requires_grad=False)
        self.sp2 = seqm_parameters['sp2']
        self.scf_converger = seqm_parameters['scf_converger']
        self.eig = seqm_parameters.get('eig', False)
        self.scf_backward = seqm_parameters.get('scf_backward', 0)
        if 'scf_backward_eps' not in seqm_parameters:
            seqm_parameters['scf_backward_eps'] = 1.0e-2
        self.scf_backward_eps = torch.nn.Parameter(torch.as_tensor(seqm_parameters['scf_backward_eps']),
                                                   requires_grad=False)
        if self.scf_backward != 0:
            raise NotImplementedError("scf_backward=1/2 is outside the B200 hot path (see DESIGN.md)")
        self._eps_cache = None

    def eps_host(self):
        """scf_eps as a python float, read back from the (possibly device-resident) parameter once"""
        key = (self.eps.data_ptr(), self.eps._version)
        if self._eps_cache is None or self._eps_cache[0] != key:
            self._eps_cache = (key, float(self.eps))
        return self._eps_cache[1]

    def forward(self, const, molsize, nHeavy, nHydro, nocc, Z, maskd, mask, atom_molid, pair_molid, idxi, idxj, ni,
                nj, xij, rij, parameters, P0=None):
        beta = torch.cat((parameters['beta_s'].unsqueeze(1), parameters['beta_p'].unsqueeze(1)), dim=1)
        Kbeta = parameters["Kbeta"] if "Kbeta" in parameters else None
        return scf_loop(const=const, molsize=molsize, nHeavy=nHeavy, nHydro=nHydro, nOccMO=nocc, maskd=maskd,
                        mask=mask, atom_molid=atom_molid, pair_molid=pair_molid, idxi=idxi, idxj=idxj, ni=ni, nj=nj,
                        xij=xij, rij=rij, Z=Z, zetas=parameters['zeta_s'], zetap=parameters['zeta_p'],
                        uss=parameters['U_ss'], upp=parameters['U_pp'], gss=parameters['g_ss'],
                        gsp=parameters['g_sp'], gpp=parameters['g_pp'], gp2=parameters['g_p2'],
                        hsp=parameters['h_sp'], beta=beta, Kbeta=Kbeta, eps=self.eps, P=P0, sp2=self.sp2,
                        scf_converger=self.scf_converger, eig=self.eig, scf_backward=self.scf_backward,
                        scf_backward_eps=self.scf_backward_eps)


class Energy(torch.nn.Module):
    """Energies (reference basics.py:253-346).  Parser -> parameters -> pair pass + SCF kernels; Eelec and
    EnucAB come out of one autograd node whose backward is the fused adjoint kernel."""

    def __init__(self, seqm_parameters):
        super().__init__()
        self.seqm_parameters = seqm_parameters
        self.method = seqm_parameters['method']
        self.parser = Parser(seqm_parameters)
        self.packpar = Pack_Parameters(seqm_parameters)
        self.hamiltonian = Hamiltonian(seqm_parameters)
        self.Hf_flag = seqm_parameters.get("Hf_flag", True)
        self._defer_status = False      # Force sets this: the SCF status is then read after the backward pass

    def forward(self, const, coordinates, species, learned_parameters=dict(), all_terms=False, P0=None, step=0,
                *args, **kwargs):
        prep = _engine.prepared_for(const, self.parser, self.packpar, species, coordinates, self.method,
                                    learned_parameters, kwargs)
        H = self.hamiltonian
        if prep is not None:
            # species-only data cached; only the pair geometry depends on the coordinates
            nmol, molsize, nHeavy, nHydro, nocc, Z, maskd = prep.nmol, prep.molsize, prep.nHeavy, prep.nHydro, \
                prep.nocc, prep.Z, prep.maskd
            atom_molid, pair_molid, idxi, idxj, parameters = prep.atom_molid, prep.pair_molid, prep.idxi, prep.idxj, \
                prep.parameters
            xij, rij = _ops.PairGeometry.apply(coordinates, prep.topo, prep.lcf)
            st = _engine.build(const, nmol, molsize, nHeavy, nHydro, nocc, Z, maskd, idxi, idxj, xij, rij, parameters,
                               self.method, run_scf=True, P0=P0, eps=H.eps_host(), sp2=H.sp2,
                               scf_converger=H.scf_converger, max_iter=_scf_loop_mod.MAX_ITER,
                               debug=_scf_loop_mod.debug, topo=prep.topo, atompar=prep.atompar, P_guess=prep.P_guess,
                               defer_status=self._defer_status)
        else:
            nmol, molsize, nHeavy, nHydro, nocc, Z, maskd, atom_molid, mask, pair_molid, ni, nj, idxi, idxj, xij, rij = \
                self.parser(const, species, coordinates, *args, **kwargs)
            if callable(learned_parameters):
                parameters = self.packpar(Z, learned_params=learned_parameters(species, coordinates))
            else:
                parameters = self.packpar(Z, learned_params=learned_parameters)
            for k, v in parameters.items():
                if torch.is_tensor(v) and v.requires_grad:
                    raise NotImplementedError("gradients with respect to parameter '%s' are outside the B200 hot path"
                                              % k)
            lt = self.parser.last_topology
            st = _engine.build(const, nmol, molsize, nHeavy, nHydro, nocc, Z, maskd, idxi, idxj, xij, rij, parameters,
                               self.method, run_scf=True, P0=P0, eps=H.eps_host(), sp2=H.sp2,
                               scf_converger=H.scf_converger, max_iter=_scf_loop_mod.MAX_ITER,
                               debug=_scf_loop_mod.debug, defer_status=self._defer_status,
                               topo=lt[1] if (lt is not None and lt[0] is idxi) else None)
        P, notconverged = st['P'], st['notconverged']
        Eelec, EnucAB = _ops.PairEnergy.apply(xij, rij, st)
        e = charge = None
        if H.eig:
            from .seqm_functions.diag import sym_eig_trunc
            from .seqm_functions.scf_loop import _charges
            e, v = sym_eig_trunc(st['F'], nHeavy, nHydro, nocc, eig_only=True)
            charge = _charges(v, nHeavy, nHydro, nmol, molsize)
        self.last_state = st
        if all_terms:
            Etot, Enuc = total_energy(nmol, pair_molid, EnucAB, Eelec)
            Eiso = prep.Eiso if prep is not None else elec_energy_isolated_atom(
                const, Z, uss=parameters['U_ss'], upp=parameters['U_pp'], gss=parameters['g_ss'],
                gpp=parameters['g_pp'], gsp=parameters['g_sp'], gp2=parameters['g_p2'], hsp=parameters['h_sp'])
            Hf, Eiso_sum = heat_formation(const, nmol, atom_molid, Z, Etot, Eiso, flag=self.Hf_flag)
            return Hf, Etot, Eelec, Enuc, Eiso_sum, EnucAB, e, P, charge, notconverged
        return Eelec, EnucAB, P, notconverged


class Force(torch.nn.Module):
    """force = -dHf/dR (reference basics.py:348-390)"""

    def __init__(self, seqm_parameters):
        super().__init__()
        self.energy = Energy(seqm_parameters)
        self.create_graph = seqm_parameters.get("2nd_grad", False)
        if self.create_graph:
            raise NotImplementedError("2nd_grad (second derivatives) is outside the B200 hot path")
        self.seqm_parameters = seqm_parameters
        # pyseqm_b200 extension (not a reference key): report SCF non-convergence of call k (the warning and
        # last_state['iterations']) when call k+1 starts instead of blocking the host at the end of call k, so the
        # host can enqueue the next step while the GPU still runs this one.  `notconverged` is returned as usual.
        self.defer_scf_warning = bool(seqm_parameters.get('defer_scf_warning', False))
        self._pending = []

    def flush_status(self):
        """resolve the deferred SCF status of earlier calls (defer_scf_warning)"""
        while self._pending:
            _engine.resolve_status(self._pending.pop(0))

    def forward(self, const, coordinates, species, learned_parameters=dict(), P0=None, step=0, *args, **kwargs):
        while self._pending and (len(self._pending) > 2 or _engine.resolve_status(self._pending[0], wait=False)):
            _engine.resolve_status(self._pending.pop(0))
        coordinates.requires_grad_(True)
        self.energy._defer_status = True
        try:
            Hf, Etot, Eelec, Enuc, Eiso, EnucAB, e, P, charge, notconverged = \
                self.energy(const, coordinates, species, learned_parameters=learned_parameters, all_terms=True, P0=P0,
                            step=step, *args, **kwargs)
        finally:
            self.energy._defer_status = False
        L = Hf.sum()
        t0 = _engine._tic(const)
        L.backward()
        _engine._toc(const, "Force", t0)
        if self.defer_scf_warning and 'status' in self.energy.last_state:
            self._pending.append(self.energy.last_state)
        else:
            _engine.resolve_status(self.energy.last_state)      # SCF status (non-convergence warning) without a mid-call sync
        force = -coordinates.grad.detach()
        coordinates.grad = torch.zeros_like(coordinates.grad)
        return force, P, Etot, Hf, Eelec, Enuc, Eiso, EnucAB, e, charge, notconverged


class GraphedForce:
    """pyseqm_b200 extension (no counterpart in the reference): ONE `Force` call over a fixed batch shape captured in a
    CUDA graph and replayed.  A Force call on the cached-species path is a fixed sequence of library launches plus a
    handful of small ATen kernels (energy sums, the autograd bookkeeping) with no host synchronisation inside, so the
    whole call - pair geometry, pair pass, one-launch SCF, adjoint pass, geometry adjoint, energies - replays as one
    graph launch: the host cost of a step drops from ~40 Python-level launches to one.

        gf = GraphedForce(seqm_parameters).to(device)
        force, P, Etot, Hf, Eelec, Enuc, Eiso, EnucAB, e, charge, notconverged = gf(const, coordinates, species)

    Same return tuple as `Force`.  The returned tensors are the graph's static output buffers: the NEXT call overwrites
    them (clone what must survive).  The graph is (re)captured when the species tensor, the batch shape or the device
    changes.  SCF non-convergence is reported through `notconverged` only (no host-side warning: nothing in a replay
    runs on the host).  Requires the cached-species path (`pair_outer_cutoff` beyond every distance, no per-call
    parameters), like the fused XL-BOMD step."""

    def __init__(self, seqm_parameters, warmup=3):
        par = dict(seqm_parameters)
        par['defer_scf_warning'] = True
        self.force = Force(par)
        self.warmup = int(warmup)
        self._key = None
        self._graph = None
        self._x = None
        self._out = None

    def to(self, device):
        self.force.to(device)
        return self

    @property
    def energy(self):
        return self.force.energy

    def _capture(self, const, coordinates, species):
        x = coordinates.detach().clone()
        side = torch.cuda.Stream(device=x.device)
        side.wait_stream(torch.cuda.current_stream(x.device))
        with torch.cuda.stream(side):
            for _ in range(self.warmup):      # allocations, module loads, the species cache, the pinned status slot
                x.grad = None
                self.force(const, x, species)
            self.force.flush_status()
        torch.cuda.current_stream(x.device).wait_stream(side)
        torch.cuda.synchronize(x.device)
        graph = torch.cuda.CUDAGraph()
        x.grad = None
        with torch.cuda.graph(graph):
            out = self.force(const, x, species)
        self.force._pending.clear()      # the status of a replay is not read on the host
        self._graph, self._x, self._out = graph, x, out

    def __call__(self, const, coordinates, species):
        key = (species.data_ptr(), species._version, tuple(coordinates.shape), str(coordinates.device))
        if key != self._key:
            if not coordinates.is_cuda:
                raise ValueError("GraphedForce needs CUDA tensors")
            self._capture(const, coordinates, species)
            self._key = key
            self._species = species      # keeps the pointer of the key alive
        self._x.data.copy_(coordinates.detach(), non_blocking=True)
        self._graph.replay()
        return self._out
