"""Element tables and unit constants (mirror of the reference's seqm/seqm_functions/constants.py:1-151;
the MOPAC7 values ev = 27.21 and a0 = 0.529167 are part of the numerical contract)."""
import torch

ev = 27.21
a0 = 0.529167
ev_kcalpmol = 23.061
overlap_cutoff = 40.0      # bohr; pairs beyond it get a zero overlap block (hcore.py:56)

_LABEL = ['0', 'H', 'He', 'Li', 'Be', ' B', ' C', ' N', ' O', ' F', 'Ne',
          'Na', 'Mg', 'Al', 'Si', ' P', ' S', 'Cl', 'Ar']
#            0   H  He  Li  Be   B   C   N   O   F  Ne  Na  Mg  Al  Si   P   S  Cl  Ar
_TABLES = {
    'tore': [0., 1, 0, 1, 2, 3, 4, 5, 6, 7, 0, 1, 2, 3, 4, 5, 6, 7, 0],
    'qn':   [0., 1, 0, 2, 2, 2, 2, 2, 2, 2, 0, 3, 3, 3, 3, 3, 3, 3, 0],
    'ussc': [0., 1, 0, 1, 2, 2, 2, 2, 2, 2, 0, 1, 2, 2, 2, 2, 2, 2, 0],
    'uppc': [0., 0, 0, 0, 0, 1, 2, 3, 4, 5, 6, 0, 0, 1, 2, 3, 4, 5, 6],
    'gssc': [0., 0, 0, 0, 1, 1, 1, 1, 1, 1, 0, 0, 1, 1, 1, 1, 1, 1, 0],
    'gspc': [0., 0, 0, 0, 0, 2, 4, 6, 8, 10, 0, 0, 0, 2, 4, 6, 8, 10, 0],
    'hspc': [0., 0, 0, 0, 0, -1, -2, -3, -4, -5, 0, 0, 0, -1, -2, -3, -4, -5, 0],
    'gp2c': [0., 0, 0, 0, 0, 0, 1.5, 4.5, 6.5, 10, 0, 0, 0, 0, 1.5, 4.5, 6.5, 10, 0],
    'gppc': [0., 0, 0, 0, 0, 0, -0.5, -1.5, -0.5, 0, 0, 0, 0, 0, -0.5, -1.5, -0.5, 0, 0],
    'eheat_kcal': [0.0, 52.102, 0.0, 38.410, 76.960, 135.700, 170.890, 113.000, 59.559, 18.890, 0.0,
                   25.850, 35.000, 79.490, 108.390, 75.570, 66.400, 28.990, 0.0],
    'mass': [0.0, 1.00790, 4.00260, 6.94000, 9.01218, 10.81000, 12.01100, 14.00670, 15.99940, 18.99840,
             20.17900, 22.98977, 24.30500, 26.98154, 28.08550, 30.97376, 32.06000, 35.45300, 39.94800],
}


class Constants(torch.nn.Module):
    """Same attributes as the reference class: tore, qn, qn_int, ussc..gppc, eheat (eV), mass, label,
    length_conversion_factor, do_timing / timing."""

    TIMING_KEYS = ("Hcore + STO Integrals", "SCF", "Force", "MD", "D*")

    def __init__(self, length_conversion_factor=(1.0 / a0), energy_conversion_factor=1.0, do_timing=False):
        super().__init__()
        self.length_conversion_factor = length_conversion_factor
        self.energy_conversion_factor = energy_conversion_factor
        self.label = list(_LABEL)
        P = lambda v: torch.nn.Parameter(torch.as_tensor(v), requires_grad=False)
        for k in ('tore', 'qn', 'ussc', 'uppc', 'gssc', 'gspc', 'hspc', 'gp2c', 'gppc', 'mass'):
            setattr(self, k, P(_TABLES[k]))
        self.qn_int = P([int(x) for x in _TABLES['qn']])
        self.eheat = P(torch.as_tensor(_TABLES['eheat_kcal']) / ev_kcalpmol)
        self.do_timing = do_timing
        # the reference only creates the dict when do_timing is set at construction (constants.py:133-140);
        # it is always present here so that callers who flip do_timing later do not crash
        self.timing = {k: [] for k in self.TIMING_KEYS}

    def forward(self):
        pass
