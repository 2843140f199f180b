"""MD and optimisation drivers with the reference's class and method names (seqm/MolecularDynamics.py).
They are host-side loops around `Force` (one kernel-backed SCF + adjoint per step); the per-step arithmetic on
coordinates / velocities is a handful of elementwise ops on device tensors."""
import torch

from . import _engine
from .basics import Force

ACC_SCALE = 0.009648532800137615            # eV/A / (g/mol) -> A/fs^2
VEL_SCALE = 0.9118367323190634e-3           # sqrt(K / amu) -> A/fs
KINETIC_ENERGY_SCALE = 1.0364270099032438e2  # amu (A/fs)^2 -> eV
EV_TO_KELVIN = 1.160451812e4


class Geometry_Optimization_SD_LS(torch.nn.Module):
    """steepest descent with a per-molecule line search over five trial step lengths (0.5 ... 1.5 times the current
    one, the lowest total energy wins; reference MolecularDynamics.py:5-87, marked "not finished" there)"""

    TRIAL = (0.5, 0.75, 1.0, 1.25, 1.5)

    def __init__(self, seqm_parameters, alpha=0.01, force_tol=1.0e-4, energy_tol=1.0e-4, max_evl=1000):
        super().__init__()
        self.seqm_parameters = seqm_parameters
        self.alpha, self.force_tol, self.energy_tol, self.max_evl = alpha, force_tol, energy_tol, max_evl
        self.force = Force(seqm_parameters)

    def onestep(self, const, coordinates, species, alpha, learned_parameters=dict(), P0=None, *args, **kwargs):
        force, P, L = self.force(const, coordinates, species, learned_parameters=learned_parameters, P0=P0,
                                 *args, **kwargs)[:3]
        P = P.detach()
        trial = torch.tensor(self.TRIAL, dtype=coordinates.dtype, device=coordinates.device)
        alphas = alpha.unsqueeze(1) * trial.reshape(1, -1)
        with torch.no_grad():
            eng = torch.zeros_like(alphas)
            for k in range(alphas.shape[1]):
                moved = (coordinates + force * alphas[:, k].reshape(-1, 1, 1)).detach()
                eng[:, k] = self.force.energy(const, moved, species, learned_parameters=learned_parameters,
                                              all_terms=True, P0=P0, *args, **kwargs)[1]
            alpha = alphas.gather(1, torch.argmin(eng, dim=1, keepdim=True)).reshape(-1)
            coordinates.add_(alpha.reshape(-1, 1, 1) * force)
        return coordinates, force, P, L, alpha

    def run(self, const, coordinates, species, learned_parameters=dict(), P=None, log=True, *args, **kwargs):
        nmol = coordinates.shape[0]
        coordinates.requires_grad_(True)
        alpha = torch.full((nmol,), float(self.alpha), dtype=coordinates.dtype, device=coordinates.device)
        Lold = torch.zeros(nmol, dtype=coordinates.dtype, device=coordinates.device)
        force_err = energy_err = None
        converged = False
        for i in range(self.max_evl):
            coordinates, force, P, Lnew, alpha = self.onestep(const, coordinates, species, alpha,
                                                              learned_parameters=learned_parameters, P0=P, *args, **kwargs)
            with torch.no_grad():
                if torch.is_tensor(coordinates.grad):
                    coordinates.grad.zero_()
                alpha = alpha.clamp_min(1.0e-3)
                force_err = torch.max(torch.abs(force))
                energy_err = (Lnew - Lold).sum() / nmol
            if log:
                print("%d %e " % (i + 1, force_err.item()), end="")
                print(" ".join("%e %e %e" % (a, b, a - b) for a, b in zip(Lnew.tolist(), Lold.tolist())))
            if force_err > self.force_tol:
                Lold = Lnew.detach().clone()
                continue
            converged = True
            break
        if not converged:
            print('not converged within %d step' % self.max_evl)
        elif log:
            print("converged with %d step, Max Force = %e (eV/Ang), dE = %e (eV)" % (i + 1, force_err.item(),
                                                                                     energy_err.item()))
        return coordinates, force_err, energy_err


class Geometry_Optimization_SD(torch.nn.Module):
    """steepest descent: x <- x + alpha * force until max |force| < force_tol (reference :103-157)"""

    def __init__(self, seqm_parameters, alpha=0.01, force_tol=1.0e-4, max_evl=1000):
        super().__init__()
        self.seqm_parameters = seqm_parameters
        self.alpha, self.force_tol, self.max_evl = alpha, force_tol, max_evl
        self.force = Force(seqm_parameters)

    def onestep(self, const, coordinates, species, learned_parameters=dict(), P0=None, *args, **kwargs):
        force, P, L = self.force(const, coordinates, species, learned_parameters=learned_parameters, P0=P0,
                                 *args, **kwargs)[:3]
        with torch.no_grad():
            coordinates.add_(self.alpha * force)
        return coordinates, force, P.detach(), L

    def run(self, const, coordinates, species, learned_parameters=dict(), P=None, log=True, *args, **kwargs):
        nmol = coordinates.shape[0]
        coordinates.requires_grad_(True)
        Lold = torch.zeros(nmol, dtype=coordinates.dtype, device=coordinates.device)
        force_err = energy_err = None
        converged = False
        for i in range(self.max_evl):
            coordinates, force, P, Lnew = self.onestep(const, coordinates, species,
                                                       learned_parameters=learned_parameters, P0=P, *args, **kwargs)
            with torch.no_grad():
                force_err = torch.max(torch.abs(force))
                energy_err = (Lnew - Lold).sum() / nmol
            if log:
                print("%d %e " % (i + 1, force_err.item()), end="")
                print(" ".join("%e" % x for x in Lnew.tolist()))
            if force_err > self.force_tol:
                Lold = Lnew.detach().clone()
                continue
            converged = True
            break
        if not converged:
            print('not converged within %d step' % self.max_evl)
        elif log:
            print("converged with %d step, Max Force = %e (eV/Ang), dE = %e (eV)" % (i + 1, force_err.item(),
                                                                                     energy_err.item()))
        return coordinates, force_err, energy_err


class Molecular_Dynamics_Basic(torch.nn.Module):
    """velocity Verlet, NVE (reference :159-393)"""

    def __init__(self, seqm_parameters, timestep=1.0, output={'molid': [0], 'thermo': 1, 'dump': 10, 'prefix': 'md'},
                 *args, **kwargs):
        super().__init__(*args, **kwargs)
        self.seqm_parameters = seqm_parameters
        self.timestep = timestep
        self.conservative_force = Force(seqm_parameters)
        self.acc_scale = ACC_SCALE
        self.vel_scale = VEL_SCALE
        self.kinetic_energy_scale = KINETIC_ENERGY_SCALE
        self.output = output

    @staticmethod
    def _masses(const, species):
        m = const.mass.detach().clone()
        m[0] = 1.0          # padding atoms: avoid 0/0 (the reference mutates const.mass in place)
        return m[species].unsqueeze(2)

    def initialize_velocity(self, const, coordinates, species, Temp=300.0, vel_com=True):
        mass = self._masses(const, species)
        scale = torch.sqrt(Temp / mass) * self.vel_scale
        velocities = torch.randn(coordinates.shape, dtype=coordinates.dtype, device=coordinates.device) * scale
        velocities[species == 0, :] = 0.0
        if vel_com:
            self.zero_com(const, species, coordinates, velocities)
        return velocities

    def zero_com(self, const, species, coordinates, velocities):
        """remove centre-of-mass position/velocity and the angular momentum, rescale to the old temperature"""
        mass = const.mass[species].unsqueeze(2)
        Mtot = torch.sum(mass, dim=1, keepdim=True)
        _, T0 = self.kinetic_energy(const, mass, species, velocities)
        with torch.no_grad():
            coordinates.sub_(torch.sum(mass * coordinates, dim=1, keepdim=True) / Mtot)
            velocities.sub_(torch.sum(mass * velocities, dim=1, keepdim=True) / Mtot)
            Lang = torch.sum(mass * torch.cross(coordinates, velocities, dim=2), dim=1)
            r2 = torch.sum(mass * torch.norm(coordinates, dim=2, keepdim=True) ** 2, dim=1, keepdim=True)
            eye = torch.eye(3, dtype=coordinates.dtype, device=coordinates.device).reshape(1, 3, 3)
            inertia = r2 * eye - torch.sum(mass.unsqueeze(3) * coordinates.unsqueeze(3) * coordinates.unsqueeze(2), dim=1)
            omega = torch.linalg.solve(inertia, Lang.unsqueeze(2))
            velocities.add_(torch.cross(coordinates, omega.reshape(-1, 1, 3).repeat(1, coordinates.shape[1], 1), dim=2))
            _, T1 = self.kinetic_energy(const, mass, species, velocities)
            velocities.mul_(torch.sqrt(T0 / T1).reshape(-1, 1, 1))

    def kinetic_energy(self, const, mass, species, velocities):
        Ek = torch.sum(0.5 * mass * velocities ** 2, dim=(1, 2)) * self.kinetic_energy_scale
        Temp = Ek * EV_TO_KELVIN / (1.5 * torch.sum(species > 0, dim=1).type(Ek.dtype))
        return Ek, Temp

    def get_force(self, const, mass, coordinates, velocities, species, learned_parameters=dict(), P0=None, step=0,
                  *args, **kwargs):
        F, P, Etot, Hf = self.conservative_force(const, coordinates, species, learned_parameters=learned_parameters,
                                                 P0=P0, step=step, *args, **kwargs)[:4]
        return F, P.detach(), Hf

    def one_step(self, const, mass, coordinates, velocities, species, acc=None, learned_parameters=dict(), P=None,
                 step=0, *args, **kwargs):
        dt = self.timestep
        t0 = _engine._tic(const)
        if not torch.is_tensor(acc):
            force, P, _ = self.get_force(const, mass, coordinates, velocities, species,
                                         learned_parameters=learned_parameters, P0=P, step=step, *args, **kwargs)
            with torch.no_grad():
                acc = force / mass * self.acc_scale
        with torch.no_grad():
            velocities.add_(0.5 * acc * dt)
            coordinates.add_(velocities * dt)
        force, P, L = self.get_force(const, mass, coordinates, velocities, species,
                                     learned_parameters=learned_parameters, P0=P, step=step, *args, **kwargs)
        with torch.no_grad():
            acc = force / mass * self.acc_scale
            velocities.add_(0.5 * acc * dt)
        _engine._toc(const, "MD", t0)
        return coordinates, velocities, acc, P, L, force

    @staticmethod
    def atomic_charges(P, n_orbital=4):
        nmol, natom = P.shape[0], P.shape[1] // n_orbital
        return P.diagonal(dim1=1, dim2=2).reshape(nmol, natom, n_orbital).sum(axis=2)

    @staticmethod
    def dipole(q, coordinates):
        return torch.sum(q.unsqueeze(2) * coordinates, axis=1)

    def screen_output(self, i, T, Ek, L, d):
        if i == 0:
            print("Step, Temp, E(kinetic), E(potential), E(total), dipole(x,y,z)")
        if (i + 1) % self.output['thermo'] == 0:
            print("%6d" % (i + 1), end="")
            for mol in self.output['molid']:
                print(" %8.2f %e %e %e %e %e %e" % (T[mol], Ek[mol], L[mol], L[mol] + Ek[mol], d[mol, 0], d[mol, 1],
                                                   d[mol, 2]), end="")
            print()

    def dump(self, i, const, species, coordinates, velocities, q, T, Ek, L, forces):
        if (i + 1) % self.output['dump'] != 0:
            return
        for mol in self.output['molid']:
            sp = species[mol].tolist()
            x, v, f, qq = coordinates[mol].tolist(), velocities[mol].tolist(), forces[mol].tolist(), q[mol].tolist()
            with open(self.output['prefix'] + "." + str(mol) + ".xyz", 'a+') as fh:
                fh.write("%d\nstep: %d, T=%6.3fK, Ek=%23.16e, Ep=%23.16e\n" % (sum(1 for z in sp if z > 0), i + 1,
                                                                              T[mol], Ek[mol], L[mol]))
                for a, z in enumerate(sp):
                    if z > 0:
                        fh.write("%2s " % const.label[z] + " ".join("%23.16e" % t for t in (x[a] + v[a] + f[a] + [qq[a]]))
                                 + "\n")

    def scale_velocities(self, i, velocities, T, scale_vel):
        if (i + 1) % scale_vel[0] == 0:
            velocities.mul_(torch.sqrt(scale_vel[1] / T).reshape(-1, 1, 1))
            return True
        return False

    def control_shift(self, velocities, Ek, Eshift):
        alpha = torch.sqrt((Ek - Eshift) / Ek)
        alpha[~torch.isfinite(alpha)] = 0.0
        velocities.mul_(alpha.reshape(-1, 1, 1))

    def _thermo(self, i, const, mass, species, coordinates, velocities, P, L, forces, q0, E0, kwargs):
        """per-step bookkeeping shared with XL_BOMD.run (reference :366-388)"""
        q = q0 - self.atomic_charges(P)
        d = self.dipole(q, coordinates)
        Ek, T = self.kinetic_energy(const, mass, species, velocities)
        if not torch.is_tensor(E0):
            E0 = L + Ek
        if 'scale_vel' in kwargs and 'control_energy_shift' in kwargs:
            raise ValueError("Can't scale velocities to fix temperature and fix energy shift at same time")
        if 'scale_vel' in kwargs and self.scale_velocities(i, velocities, T, kwargs["scale_vel"]):
            Ek, T = self.kinetic_energy(const, mass, species, velocities)
        if kwargs.get('control_energy_shift'):
            self.control_shift(velocities, Ek, Ek + L - E0)
            Ek, T = self.kinetic_energy(const, mass, species, velocities)
        self.screen_output(i, T, Ek, L, d)
        self.dump(i, const, species, coordinates, velocities, q, T, Ek, L, forces)
        return E0

    def run(self, const, steps, coordinates, velocities, species, learned_parameters=dict(), reuse_P=True,
            remove_com=[False, 1000], *args, **kwargs):
        mass = self._masses(const, species)
        acc = P = E0 = None
        q0 = const.tore[species]
        fkw = {k: v for k, v in kwargs.items() if k not in ('scale_vel', 'control_energy_shift')}
        for i in range(steps):
            coordinates, velocities, acc, P, L, forces = self.one_step(
                const, mass, coordinates, velocities, species, acc=acc, learned_parameters=learned_parameters, P=P,
                step=i, *args, **fkw)
            with torch.no_grad():
                Pq = P
                if not reuse_P:
                    P = None
                if remove_com[0] and i % remove_com[1] == 0:
                    self.zero_com(const, species, coordinates, velocities)
                E0 = self._thermo(i, const, mass, species, coordinates, velocities, Pq, L, forces, q0, E0, kwargs)
        return coordinates, velocities, acc


class Molecular_Dynamics_Langevin(Molecular_Dynamics_Basic):
    """Langevin thermostat: F = Fc - m v / damp + sqrt(2 kB T m / (dt damp)) R(t)  (reference :396-436).
    As in the reference the parent is constructed with timestep=1.0 whatever `timestep` is passed."""

    def __init__(self, seqm_parameters, timestep=1.0, damp=1.0, Temp=300.0,
                 output={'molid': [0], 'thermo': 1, 'dump': 10, 'prefix': 'md'}, *args, **kwargs):
        super().__init__(seqm_parameters, timestep=1.0, output=output, *args, **kwargs)
        self.damp = damp
        self.T = Temp
        self.Fr_scale = 0.09450522179973914

    def get_force(self, const, masses, coordinates, velocities, species, learned_parameters=dict(), P0=None, step=0,
                  *args, **kwargs):
        Fc, P, L = self.conservative_force(const, coordinates, species, learned_parameters=learned_parameters,
                                           P0=P0)[:3]
        Ff = -masses * velocities / self.damp / self.acc_scale
        Fr = self.Fr_scale * torch.sqrt(2.0 * self.T * masses / self.timestep / self.damp) * \
            torch.randn(Fc.shape, dtype=Fc.dtype, device=Fc.device)
        F = Fc + Ff + Fr
        F[species == 0, :] = 0.0
        return F, P.detach(), L
