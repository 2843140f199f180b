"""A user script written against the REFERENCE's import names (`import seqm`, `from seqm.basics import Energy` ...):
energy terms of a batch of homonuclear diatomics with explicit charges, orbital energies and per-orbital charges, in
the configuration of the reference's tests/test.py (PM3, adaptive mixing, finite pair cutoff, Hf_flag off, eig on).
tools/make_golden_dropin.py runs THIS file against the unmodified reference (CPU) and stores what it returns;
tests/test_gpu_dropin.py runs the same file with `sys.modules['seqm'] = pyseqm_b200` on the GPU and compares."""
import torch


def main(device, check_degeneracy=False, dx=1.0):
    import seqm
    from seqm.basics import Energy
    from seqm.seqm_functions.constants import Constants
    seqm.seqm_functions.scf_loop.debug = False
    seqm.seqm_functions.diag.CHECK_DEGENERACY = check_degeneracy
    torch.set_default_dtype(torch.float64)
    species = torch.as_tensor([[1, 1], [6, 6], [7, 7], [8, 8]], dtype=torch.int64, device=device)
    charges = torch.zeros(4, dtype=torch.double, device=device)
    coordinates = torch.tensor([[[0.0, 0.0, 0.0], [dx, 0.0, 0.0]]], device=device).repeat(4, 1, 1).contiguous()
    seqm_parameters = {'method': 'PM3', 'scf_eps': 1.0e-9, 'scf_converger': [1, 0.9], 'sp2': [False, 1.0e-5],
                       'elements': [0, 1, 6, 7, 8], 'learned': [], 'pair_outer_cutoff': 1.0e3, 'Hf_flag': False,
                       'eig': True}
    if hasattr(seqm, '__path__') and 'pyseqm_b200' not in seqm.__name__:
        import os
        seqm_parameters['parameter_file_dir'] = os.path.join(list(seqm.__path__)[0], 'params') + '/'
    const = Constants().to(device)
    eng = Energy(seqm_parameters).to(device)
    try:
        out = eng(const, coordinates, species, all_terms=True, charges=charges)
    finally:
        seqm.seqm_functions.diag.CHECK_DEGENERACY = False
    names = ('Etot_m_Eiso', 'Etot', 'Eelec', 'Enuc', 'Eiso', 'EnucAB', 'e', 'P', 'charge', 'notconverged')
    return {k: v.detach().cpu() for k, v in zip(names, out)}
