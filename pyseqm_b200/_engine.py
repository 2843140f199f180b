"""Host-side orchestration of one Hamiltonian build: topology (cached per pair list) -> per-atom records ->
pair pass (hcore) -> device-resident SCF.  Used by seqm_functions.scf_loop.scf_loop, basics.Energy and
XLBOMD.EnergyXL.

`Prepared` is the fast path of Energy / EnergyXL: everything that depends only on the species tensor (the
Parser's index tensors, topology, per-atom parameters and records, the initial density guess, isolated-atom
energies) is computed once and reused while the same species tensor comes back (MD, geometry scans); per call
only the pair geometry (one kernel, with a kernel adjoint) is recomputed.  It applies when the pair list cannot
change with the coordinates (pair_outer_cutoff effectively infinite) and no per-call parameters are given."""
import threading
import time
import warnings

import torch

from . import _ops
from ._topology import PairTopology

_TOPO_CACHE = []      # small LRU of (key tensors, PairTopology)
_TOPO_CACHE_SIZE = 4
_PREP_CACHE = []
_PREP_CACHE_SIZE = 4
_PREP_LOCK = threading.Lock()
_TOKENS = iter(range(1, 1 << 62))      # identity of a Pack_Parameters instance in cache keys (id() can be recycled)


def _token(obj):
    t = getattr(obj, '_seqm_b200_token', None)
    if t is None:
        with _PREP_LOCK:
            t = next(_TOKENS)
        obj._seqm_b200_token = t
    return t      # BucketedBatch may drive several Force modules from worker threads
ALL_PAIRS_CUTOFF = 1.0e6      # Angstrom: beyond any molecule, the pair list is the full a<b list


def get_topology(nmol, molsize, Z, maskd, idxi, idxj, nHeavy, nHydro, nocc):
    """PairTopology for these Parser tensors; reused while the pair list is unchanged.  Validation compares
    the index tensors on the device (one sync)."""
    for k, (cz, cm, ci, cj, co, topo) in enumerate(_TOPO_CACHE):
        if (topo.nmol == nmol and topo.molsize == molsize and cz.shape == Z.shape and ci.shape == idxi.shape
                and cz.device == Z.device):
            same = torch.equal(ci, idxi) and torch.equal(cj, idxj) and torch.equal(cz, Z) and \
                torch.equal(cm, maskd) and torch.equal(co, nocc)
            if same:
                if k:
                    _TOPO_CACHE.insert(0, _TOPO_CACHE.pop(k))
                return topo
    topo = PairTopology(nmol, molsize, Z, maskd, idxi, idxj, nHeavy, nHydro, nocc)
    _TOPO_CACHE.insert(0, (Z.clone(), maskd.clone(), idxi.clone(), idxj.clone(), nocc.clone(), topo))
    del _TOPO_CACHE[_TOPO_CACHE_SIZE:]
    return topo


def initial_density(const, topo, Z, maskd):
    """scf_loop.py:700-710: tore/4 on the four orbitals of a heavy atom, 1 on the hydrogen s orbital."""
    B, S = topo.nmol, topo.molsize
    Pb = torch.zeros(B * S * S, 4, 4, dtype=torch.float64, device=Z.device)
    heavy = Z > 1
    d = torch.zeros(Z.shape[0], 4, dtype=torch.float64, device=Z.device)
    d[heavy] = (const.tore[Z[heavy]].to(torch.float64) / 4.0).unsqueeze(1)
    d[Z == 1, 0] = 1.0
    Pb[maskd] = torch.diag_embed(d)
    return Pb.reshape(B, S, S, 4, 4).transpose(2, 3).reshape(B, 4 * S, 4 * S).contiguous()


def _tic(const):
    if const.do_timing:
        torch.cuda.synchronize()
        return time.time()
    return None


def _toc(const, key, t0):
    if const.do_timing:
        torch.cuda.synchronize()
        const.timing.setdefault(key, []).append(time.time() - t0)


def build(const, nmol, molsize, nHeavy, nHydro, nocc, Z, maskd, idxi, idxj, xij, rij, parameters, method,
          run_scf=True, P0=None, eps=1.0e-4, sp2=(False,), scf_converger=(1,), max_iter=1000, debug=False,
          topo=None, atompar=None, P_guess=None, defer_status=False):
    """-> state dict: topo, atompar, w, M, EnucAB, pv and (if run_scf) P, F, Eelec, notconverged, niter"""
    for k in ('Kbeta', 'g_ss_nuc'):
        if k in parameters:
            raise NotImplementedError("parameter '%s' (learned-parameter feature) is outside the B200 hot path" % k)
    if topo is None:
        topo = get_topology(int(nmol), int(molsize), Z, maskd, idxi, idxj, nHeavy, nHydro, nocc)
    t0 = _tic(const)
    with torch.no_grad():
        if atompar is None:
            atompar = _ops.atom_pack(topo, parameters, method)
        w, M, EnucAB, pv = _ops.hcore(topo, atompar, xij, rij, parameters['U_ss'], parameters['U_pp'], method)
    _toc(const, "Hcore + STO Integrals", t0)
    st = dict(topo=topo, atompar=atompar, w=w, M=M, EnucAB=EnucAB, pv=pv, method=method)
    par5 = tuple(parameters[k] for k in ('g_ss', 'g_pp', 'g_sp', 'g_p2', 'h_sp'))
    st['par5'] = par5
    if run_scf:
        t0 = _tic(const)
        with torch.no_grad():
            if torch.is_tensor(P0):
                P = P0
            elif P_guess is not None:
                P = P_guess
            else:
                P = initial_density(const, topo, Z, maskd)
            status = None
            if defer_status and not const.do_timing:
                status = torch.zeros(2, dtype=torch.int64).pin_memory()
            P, F, E, nc, nit, done = _ops.scf(topo, w, M, par5, P, float(eps), list(scf_converger), list(sp2), max_iter,
                                              debug, pv, status=status)
        st.update(P=P, F=F, Eelec=E, notconverged=nc, niter=nit, nmol=int(nmol), max_iter=max_iter)
        if done is None:
            st['status'] = status          # resolved by resolve_status() once the stream has been synchronised
            st['status_event'] = torch.cuda.Event()
            st['status_event'].record()
            st['iterations'] = None
        else:
            st['iterations'] = done[0]
            _warn_unconverged(done[1], nmol, max_iter)
        _toc(const, "SCF", t0)
    return st


def _warn_unconverged(left, nmol, max_iter):
    if left:
        warnings.warn("SCF for %d/%d molecules doesn't converge after %d iterations" % (left, nmol, max_iter))


def resolve_status(st, wait=True):
    """finish a build() made with defer_status: wait for the SCF's status record (not for the whole stream), read
    {left, iterations}, warn like the reference.  wait=False: only if the record has already arrived; returns whether
    the status is resolved"""
    status = st.get('status')
    if status is None:
        return True
    ev = st.get('status_event')
    if ev is not None:
        if not wait and not ev.query():
            return False
        ev.synchronize()
    else:
        torch.cuda.current_stream().synchronize()
    st.pop('status', None)
    st.pop('status_event', None)
    left, iters = int(status[0]), int(status[1])
    st['iterations'] = iters
    _warn_unconverged(left, st['nmol'], st['max_iter'])
    return True


def hcore_matrix(M, nmol, molsize):
    return M.reshape(nmol, molsize, molsize, 4, 4).transpose(2, 3).reshape(nmol, 4 * molsize, 4 * molsize)


class Prepared:
    """species-only data of one batch (see module docstring)"""

    def __init__(self, const, parser, packpar, species, coordinates, method):
        from .seqm_functions.energy import elec_energy_isolated_atom
        with torch.no_grad():
            out = parser(const, species, coordinates.detach())
        (self.nmol, self.molsize, self.nHeavy, self.nHydro, self.nocc, self.Z, self.maskd, self.atom_molid, self.mask,
         self.pair_molid, self.ni, self.nj, self.idxi, self.idxj, _, _) = out
        self.species = species
        self.key = (species.data_ptr(), species._version, tuple(species.shape), str(species.device), method,
                    _token(packpar))
        self.parameters = packpar(self.Z, learned_params=dict())
        lt = getattr(parser, 'last_topology', None)
        if lt is not None and lt[0] is self.idxi:          # tables already built on the device by the Parser kernels
            self.topo = lt[1]
        else:
            self.topo = PairTopology(self.nmol, self.molsize, self.Z, self.maskd, self.idxi, self.idxj, self.nHeavy,
                                     self.nHydro, self.nocc)
        self.atompar = _ops.atom_pack(self.topo, self.parameters, method)
        self.P_guess = initial_density(const, self.topo, self.Z, self.maskd)
        p = self.parameters
        self.Eiso = elec_energy_isolated_atom(const, self.Z, uss=p['U_ss'], upp=p['U_pp'], gss=p['g_ss'],
                                              gpp=p['g_pp'], gsp=p['g_sp'], gp2=p['g_p2'], hsp=p['h_sp'])
        self.lcf = float(const.length_conversion_factor)


def prepared_for(const, parser, packpar, species, coordinates, method, learned_parameters, kwargs):
    """Prepared object if the fast path applies, else None"""
    if kwargs or callable(learned_parameters) or len(learned_parameters) > 0:
        return None
    if not (parser.outercutoff >= ALL_PAIRS_CUTOFF) or not species.is_cuda:
        return None
    key = (species.data_ptr(), species._version, tuple(species.shape), str(species.device), method, _token(packpar))
    with _PREP_LOCK:
        for k, prep in enumerate(_PREP_CACHE):
            if prep.key == key:
                if k:
                    _PREP_CACHE.insert(0, _PREP_CACHE.pop(k))
                return prep
    prep = Prepared(const, parser, packpar, species, coordinates, method)
    with _PREP_LOCK:
        _PREP_CACHE.insert(0, prep)
        del _PREP_CACHE[_PREP_CACHE_SIZE:]
    return prep
