"""A user script against the reference's import names: forces and heats of formation of a small padded batch
(formaldehyde-class molecules, AM1, Pulay converger), the shape of the reference's tests/test2.py / test8.py."""
import torch


def main(device):
    import seqm
    from seqm.basics import Force
    from seqm.seqm_functions.constants import Constants
    seqm.seqm_functions.scf_loop.debug = False
    torch.set_default_dtype(torch.float64)
    species = torch.as_tensor([[8, 6, 1, 1], [8, 6, 1, 1], [8, 8, 6, 0]], dtype=torch.int64, device=device)
    coordinates = torch.tensor([[[0.00, 0.00, 0.00], [1.22, 0.00, 0.00], [1.82, 0.94, 0.00], [1.82, -0.94, 0.00]],
                                [[0.00, 0.00, 0.00], [1.22, 0.00, 0.00], [1.82, 0.94, 0.00], [1.82, -0.94, 0.00]],
                                [[0.00, 0.00, 0.00], [1.23, 0.00, 0.00], [1.82, 0.94, 0.00], [0.00, 0.00, 0.00]]],
                               device=device)
    coordinates[1, :, 2] += torch.tensor([0.05, -0.03, 0.02, 0.04], device=device)
    seqm_parameters = {'method': 'AM1', 'scf_eps': 1.0e-8, 'scf_converger': [2, 0.0], 'sp2': [False, 1.0e-5],
                       'elements': [0, 1, 6, 8], 'learned': [], 'pair_outer_cutoff': 1.0e10}
    if hasattr(seqm, '__path__') and 'pyseqm_b200' not in seqm.__name__:
        import os
        seqm_parameters['parameter_file_dir'] = os.path.join(list(seqm.__path__)[0], 'params') + '/'
    const = Constants().to(device)
    force, P, Etot, Hf, Eelec, Enuc, Eiso, EnucAB, e, charge, notconverged = \
        Force(seqm_parameters).to(device)(const, coordinates, species)
    assert coordinates.requires_grad and float(coordinates.grad.abs().max()) == 0.0      # Force's documented side effects
    return dict(force=force.detach().cpu(), P=P.detach().cpu(), Etot=Etot.detach().cpu(), Hf=Hf.detach().cpu(),
                Eelec=Eelec.detach().cpu(), Enuc=Enuc.detach().cpu(), Eiso=Eiso.detach().cpu(),
                EnucAB=EnucAB.detach().cpu(), notconverged=notconverged.cpu())
