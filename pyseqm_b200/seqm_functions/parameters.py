"""MOPAC parameter tables (mirror of seqm/seqm_functions/parameters.py:4-31)."""
import torch


def params(method='MNDO', elements=[1, 6, 7, 8],
           parameters=['U_ss', 'U_pp', 'zeta_s', 'zeta_p', 'beta_s', 'beta_p',
                       'g_ss', 'g_sp', 'g_pp', 'g_p2', 'h_sp', 'alpha'],
           root_dir='./params/MOPAC/'):
    """p[Z, i] = value of parameters[i] for element Z (rows of elements not listed stay zero)."""
    table = torch.zeros((max(elements) + 1, len(parameters)))
    with open(root_dir + "parameters_" + method + "_MOPAC.csv") as f:
        header = f.readline().strip().replace(' ', '').split(',')
        cols = [header.index(name) for name in parameters]
        for line in f:
            t = line.strip().replace(' ', '').split(',')
            if not t or not t[0]:
                continue
            z = int(t[0])
            if z in elements:
                table[z, :] = torch.tensor([float(t[c]) for c in cols])
    return torch.nn.Parameter(table, requires_grad=False)
