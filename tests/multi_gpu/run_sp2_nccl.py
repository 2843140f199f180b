"""torchrun script (one process per GPU, NCCL): the tile-sharded SP2 of a water cluster against the single-GPU
kernel on the same Fock matrix, then a full Force call with the SCF's SP2 sharded.  Prints SP2-NCCL-OK on rank 0."""
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
torch.set_default_dtype(torch.float64)


def main():
    rank, world = int(os.environ['RANK']), int(os.environ['WORLD_SIZE'])
    local = int(os.environ.get('LOCAL_RANK', rank))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    dist.init_process_group('nccl', device_id=dev)
    from pyseqm_b200 import _ops, synth
    from pyseqm_b200.basics import Force
    from pyseqm_b200.distributed import disable_sp2_sharding, enable_sp2_sharding
    from pyseqm_b200.seqm_functions.constants import Constants
    nwater = int(os.environ.get('NWATER', '100'))
    sp, xyz = synth.water_cluster(nwater, seed=3)
    par = dict(method='PM3', elements=[0, 1, 8], scf_eps=1.0e-6, scf_converger=[2, 0.0], sp2=[True, 1.0e-5], learned=[],
               pair_outer_cutoff=1.0e10)
    species, coords = torch.as_tensor(sp).to(dev), torch.as_tensor(xyz).to(dev)
    const = Constants().to(dev)
    fm = Force(par).to(dev)
    t0 = time.time()
    out = fm(const, coords.clone(), species)
    torch.cuda.synchronize()
    t_single = time.time() - t0
    st = fm.energy.last_state
    topo, F = st['topo'], st['F']
    S = species.shape[1]
    D0, n0 = _ops.sp2_density(F, topo.nheavy, topo.nhydro, topo.nocc, S, topo.mmax, 1e-5, want_niter=True)
    comm = enable_sp2_sharding()
    try:
        for p2p in (1, 0, 1):      # peer-to-peer stores over NVLink / ncclAllGather of the records / again (buffers reused)
            _ops.lib().seqm_sp2_nccl_p2p(p2p)
            D1, n1 = _ops.sp2_density(F, topo.nheavy, topo.nhydro, topo.nocc, S, topo.mmax, 1e-5, want_niter=True)
            torch.cuda.synchronize()
            assert torch.equal(n0, n1), (p2p, n0, n1)
            assert torch.equal(D0, D1), (p2p, float((D0 - D1).abs().max()))
        # every rank holds the same matrix
        ref = D1.clone()
        dist.broadcast(ref, 0)
        assert torch.equal(ref, D1)
        t0 = time.time()
        fm2 = Force(par).to(dev)
        out2 = fm2(const, coords.clone(), species)
        torch.cuda.synchronize()
        t_sharded = time.time() - t0
        assert float((out[2] - out2[2]).abs().max()) < 1e-6 and float((out[0] - out2[0]).abs().max()) < 1e-5
        # replicated ranks must agree bit for bit on what the library computes (Etot itself goes through torch's
        # atomic segment sums, reproducible only to 1e-10)
        el = fm2.energy.last_state['Eelec'].detach().contiguous()
        e_all = [torch.empty_like(el) for _ in range(world)]
        dist.all_gather(e_all, el)
        assert all(torch.equal(e_all[0], x) for x in e_all), "ranks disagree on Eelec"
        assert torch.equal(fm2.energy.last_state['F'], st['F']), "sharded SCF differs from the single-GPU SCF"
    finally:
        disable_sp2_sharding()
    dist.barrier()
    if rank == 0:
        print("SP2-NCCL-OK world=%d waters=%d m=%d sp2_iterations=%d  Force single %.3f s  sharded %.3f s" %
              (world, nwater, int(topo.mmax), int(n1[0]), t_single, t_sharded))
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
