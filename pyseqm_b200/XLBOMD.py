"""XL-BOMD (extended-Lagrangian Born-Oppenheimer MD) with the reference's classes (seqm/XLBOMD.py:40-368):
one pair pass + one Fock build with the dynamical density P + ONE density solve (SP2 or eigh) + one adjoint
pass per MD step, no SCF loop."""
import torch
from torch.autograd import grad

from . import _engine, _ops
from .basics import Force, Pack_Parameters, Parser
from .MolecularDynamics import Molecular_Dynamics_Basic
from .seqm_functions.energy import elec_energy_isolated_atom, heat_formation, total_energy


def elec_energy_xl(D, P, F, Hcore):
    """E(D,P) = tr(D F) - 1/2 tr((F - Hcore) P)   (reference XLBOMD.py:40-52); forward only"""
    nmol, n4, _ = P.shape
    S = n4 // 4
    M = Hcore.reshape(nmol, S, 4, S, 4).transpose(2, 3).reshape(nmol * S * S, 4, 4).contiguous()
    return _ops.elec_energy(P, F, M, nmol, S, D=D)


class EnergyXL(torch.nn.Module):
    def __init__(self, seqm_parameters):
        super().__init__()
        self.method = seqm_parameters['method']
        self.seqm_parameters = seqm_parameters
        self.parser = Parser(seqm_parameters)
        self.packpar = Pack_Parameters(seqm_parameters)
        self.Hf_flag = seqm_parameters.get("Hf_flag", True)

    def forward(self, const, coordinates, species, P, learned_parameters=dict(), all_terms=False, step=0, *args,
                **kwargs):
        """P: dynamical field density; D: density from F(P) (SP2 / eigh), returned last"""
        prep = _engine.prepared_for(const, self.parser, self.packpar, species, coordinates, self.method,
                                    learned_parameters, kwargs)
        if prep is not None:
            nmol, molsize, nHeavy, nHydro, nocc, Z, maskd = prep.nmol, prep.molsize, prep.nHeavy, prep.nHydro, \
                prep.nocc, prep.Z, prep.maskd
            atom_molid, pair_molid, idxi, idxj, parameters = prep.atom_molid, prep.pair_molid, prep.idxi, prep.idxj, \
                prep.parameters
            xij, rij = _ops.PairGeometry.apply(coordinates, prep.topo, prep.lcf)
            st = _engine.build(const, nmol, molsize, nHeavy, nHydro, nocc, Z, maskd, idxi, idxj, xij, rij, parameters,
                               self.method, run_scf=False, topo=prep.topo, atompar=prep.atompar)
        else:
            nmol, molsize, nHeavy, nHydro, nocc, Z, maskd, atom_molid, mask, pair_molid, ni, nj, idxi, idxj, xij, rij = \
                self.parser(const, species, coordinates, *args, **kwargs)
            if callable(learned_parameters):
                parameters = self.packpar(Z, learned_params=learned_parameters(species, coordinates))
            else:
                parameters = self.packpar(Z, learned_params=learned_parameters)
            lt = self.parser.last_topology
            st = _engine.build(const, nmol, molsize, nHeavy, nHydro, nocc, Z, maskd, idxi, idxj, xij, rij, parameters,
                               self.method, run_scf=False, topo=lt[1] if (lt is not None and lt[0] is idxi) else None)
        topo = st['topo']
        sp2 = self.seqm_parameters['sp2']
        with torch.no_grad():
            Pd = P.detach()
            F = _ops.fock(topo, st['w'], Pd, st['M'], st['par5'], pv=st['pv'])
            t0 = _engine._tic(const)
            if sp2[0]:
                D = _ops.sp2_density(F, topo.nheavy, topo.nhydro, topo.nocc, molsize, topo.mmax, sp2[1])
            else:
                D = _ops.eigh_density(F, topo.nheavy, topo.nhydro, topo.nocc, molsize, topo.mmax)
            _engine._toc(const, "D*", t0)
            Eel = _ops.elec_energy(Pd, F, st['M'], nmol, molsize, D=D)
        st.update(P=Pd, D=D, F=F, Eelec=Eel)
        Eelec, EnucAB = _ops.PairEnergy.apply(xij, rij, st)
        if all_terms:
            Etot, Enuc = total_energy(nmol, pair_molid, EnucAB, Eelec)
            Eiso = prep.Eiso if prep is not None else elec_energy_isolated_atom(
                const, Z, uss=parameters['U_ss'], upp=parameters['U_pp'], gss=parameters['g_ss'],
                gpp=parameters['g_pp'], gsp=parameters['g_sp'], gp2=parameters['g_p2'], hsp=parameters['h_sp'])
            Hf, Eiso_sum = heat_formation(const, nmol, atom_molid, Z, Etot, Eiso, flag=self.Hf_flag)
            return Hf, Etot, Eelec, Enuc, Eiso_sum, EnucAB, D
        return Eelec, EnucAB, D


class ForceXL(torch.nn.Module):
    def __init__(self, seqm_parameters):
        super().__init__()
        self.energy = EnergyXL(seqm_parameters)
        self.seqm_parameters = seqm_parameters

    def forward(self, const, coordinates, species, P, learned_parameters=dict(), step=0, *args, **kwargs):
        coordinates.requires_grad_(True)
        Hf, Etot, Eelec, Enuc, Eiso, EnucAB, D = self.energy(const, coordinates, species, P,
                                                             learned_parameters=learned_parameters, all_terms=True,
                                                             step=step, *args, **kwargs)
        L = Hf.sum()
        t0 = _engine._tic(const)
        g = grad(L, [coordinates])[0]
        _engine._toc(const, "Force", t0)
        coordinates.grad = torch.zeros_like(g)
        return -g, Hf, D.detach()


class XL_BOMD(Molecular_Dynamics_Basic):
    """velocity Verlet with the auxiliary density propagated by the k-term dissipative scheme
    (coefficient table of reference XLBOMD.py:238-245).  As in the reference, timestep is 1.0 fs whatever is
    passed (the parent is constructed with timestep=1.0)."""

    COEFFS = {3: [1.69, 150e-3, -2.0, 3.0, 0.0, -1.0],
              4: [1.75, 57e-3, -3.0, 6.0, -2.0, -2.0, 1.0],
              5: [1.82, 18e-3, -6.0, 14.0, -8.0, -3.0, 4.0, -1.0],
              6: [1.84, 5.5e-3, -14.0, 36.0, -27.0, -2.0, 12.0, -6.0, 1.0],
              7: [1.86, 1.6e-3, -36.0, 99.0, -88.0, 11.0, 32.0, -25.0, 8.0, -1.0],
              8: [1.88, 0.44e-3, -99.0, 286.0, -286.0, 78.0, 78.0, -90.0, 42.0, -10.0, 1.0],
              9: [1.89, 0.12e-3, -286.0, 858.0, -936.0, 364.0, 168.0, -300.0, 184.0, -63.0, 12.0, -1.0]}

    def __init__(self, seqm_parameters, timestep=1.0, k=5, output={'molid': [0], 'thermo': 1, 'dump': 10, 'prefix': 'md'},
                 *args, **kwargs):
        super().__init__(seqm_parameters, timestep=1.0, output=output, *args, **kwargs)
        self.conservative_force = ForceXL(seqm_parameters)
        self.force0 = Force(seqm_parameters)
        self.coeffs = self.COEFFS
        self.k, self.m = k, k + 1
        self.kappa, self.alpha = self.COEFFS[k][0], self.COEFFS[k][1]
        cc = 1.00
        tmp = torch.as_tensor(self.COEFFS[k][2:]) * self.alpha
        self.coeff_D = cc * self.kappa
        tmp[0] += (2.0 - cc * self.kappa)
        tmp[1] -= 1.0
        self.coeff = torch.nn.Parameter(tmp.repeat(2), requires_grad=False)

    def initialize(self, const, mass, coordinates, species, learned_parameters=dict(), *args, **kwargs):
        f, D, _ = self.force0(const, coordinates, species, learned_parameters=learned_parameters, *args, **kwargs)[:3]
        with torch.no_grad():
            acc = f / mass * self.acc_scale
        return acc, D.detach()

    def get_force(self):
        pass

    def _fused_step(self, const, step, mass, coordinates, velocities, species, acc, D, P, Pt, learned_parameters, kwargs):
        """the XLStep object for this state, or None when the fused form does not apply"""
        if not self.seqm_parameters.get('xl_fused', True) or const.do_timing or not coordinates.is_cuda:
            return None
        xs = getattr(self, '_xlstep', None)
        if xs is not None and xs.matches(coordinates, velocities, acc, D, P, Pt) and xs.species is species:
            if int(step) != xs.step0 + xs.nsteps:          # the caller jumped in the step count: follow it
                xs.step_dev.fill_(int(step))
                xs.step0 = int(step) - xs.nsteps
                xs.graph = None
            return xs
        en = self.conservative_force.energy
        prep = _engine.prepared_for(const, en.parser, en.packpar, species, coordinates, en.method, learned_parameters,
                                    kwargs)
        ok = all(torch.is_tensor(t) and t.is_cuda and t.dtype == torch.float64 and t.is_contiguous()
                 for t in (coordinates, velocities, acc, D, P, Pt))
        if prep is None or not ok:
            return None
        from ._xlstep import XLStep
        xs = XLStep(self, const, prep, species, coordinates, velocities, acc, D, P, Pt, mass, int(step),
                    graph=bool(self.seqm_parameters.get('xl_cuda_graph', False)))
        xs.species, xs.step0 = species, int(step)
        self._xlstep = xs
        return xs

    def one_step(self, const, step, mass, coordinates, velocities, species, acc, D, P, Pt, learned_parameters=dict(),
                 *args, **kwargs):
        """pyseqm_b200: when the species-only data is cacheable the step runs as a fixed sequence of library launches on
        the caller's state tensors, updated IN PLACE (_xlstep.XLStep; `xl_fused=False` in seqm_parameters restores the
        autograd form below, `xl_cuda_graph=True` replays the step as one CUDA graph)."""
        xs = None if args else self._fused_step(const, step, mass, coordinates, velocities, species, acc, D, P, Pt,
                                                learned_parameters, kwargs)
        if xs is not None:
            Hf, force = xs.step()
            return coordinates, velocities, acc, D, P, Pt, Hf, force
        dt = self.timestep
        t0 = _engine._tic(const)
        with torch.no_grad():
            velocities.add_(0.5 * acc * dt)
            coordinates.add_(velocities * dt)
            cindx = step % self.m
            P = self.coeff_D * D + torch.sum(self.coeff[cindx:(cindx + self.m)].reshape(-1, 1, 1, 1) * Pt, dim=0)
            Pt[(self.m - 1 - cindx)] = P
        force, Hf, D = self.conservative_force(const, coordinates, species, P, learned_parameters=learned_parameters,
                                               step=step, *args, **kwargs)
        with torch.no_grad():
            acc = force / mass * self.acc_scale
            velocities.add_(0.5 * acc * dt)
        _engine._toc(const, "MD", t0)
        return coordinates, velocities, acc, D, P, Pt, Hf, force

    def run(self, const, steps, coordinates, velocities, species, learned_parameters=dict(), Pt=None, *args, **kwargs):
        mass = self._masses(const, species)
        fkw = {k: v for k, v in kwargs.items() if k not in ('scale_vel', 'control_energy_shift')}
        acc, D = self.initialize(const, mass, coordinates, species, learned_parameters=learned_parameters, *args, **fkw)
        with torch.no_grad():
            if not torch.is_tensor(Pt):
                Pt = D.unsqueeze(0).expand((self.m,) + D.shape).clone()
            P = D.clone()
        q0 = const.tore[species]
        E0 = None
        for i in range(steps):
            coordinates, velocities, acc, D, P, Pt, L, forces = self.one_step(
                const, i, mass, coordinates, velocities, species, acc, D, P, Pt, learned_parameters=learned_parameters,
                *args, **fkw)
            with torch.no_grad():
                E0 = self._thermo(i, const, mass, species, coordinates, velocities, P, L, forces, q0, E0, kwargs)
        return coordinates, velocities, acc, P, Pt
