"""world_size-2 gloo test of the molecule sharding / gather host logic (no GPU)."""
import os
import socket
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, nmol, q):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from pyseqm_b200.distributed import gather_molecules, shard_batch
    species = torch.arange(nmol * 3).reshape(nmol, 3)
    coords = torch.arange(nmol * 9, dtype=torch.float64).reshape(nmol, 3, 3)
    sp, xyz, (lo, hi) = shard_batch(species, coords)
    assert sp.shape[0] == hi - lo and torch.equal(xyz, coords[lo:hi])
    energy = xyz.sum(dim=(1, 2)) * 2.0                      # stands for a per-molecule result
    force = -xyz
    e_all = gather_molecules(energy, nmol)
    f_all = gather_molecules(force, nmol)
    ok = torch.equal(e_all, coords.sum(dim=(1, 2)) * 2.0) and torch.equal(f_all, -coords)
    q.put((rank, lo, hi, bool(ok)))
    dist.destroy_process_group()


def test_shard_bounds_cover_batch():
    from pyseqm_b200.distributed import shard_bounds
    for nmol in (0, 1, 7, 8, 1024, 1025):
        for world in (1, 2, 3, 8):
            b = shard_bounds(nmol, world)
            assert b[0][0] == 0 and b[-1][1] == nmol
            assert all(b[k][1] == b[k + 1][0] for k in range(world - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1


def test_shard_and_gather_world2_gloo():
    world = 2
    for nmol in (8, 7):                                    # even and ragged split
        port = _free_port()
        ctx = mp.get_context('spawn')
        q = ctx.Queue()
        procs = [ctx.Process(target=_worker, args=(r, world, port, nmol, q)) for r in range(world)]
        for p in procs:
            p.start()
        res = sorted(q.get(timeout=120) for _ in range(world))
        for p in procs:
            p.join(timeout=60)
            assert p.exitcode == 0
        assert all(r[3] for r in res)
        assert res[0][1] == 0 and res[0][2] == res[1][1] and res[1][2] == nmol


def test_sp2_tile_plan_partitions_the_upper_triangle():
    """seqm_sp2_shard_plan (host part of the tile-sharded SP2 of config 5): contiguous ranges that cover
    [0, ntile) exactly once, sized in whole sub-chunks"""
    import ctypes
    from pyseqm_b200 import _lib
    from pyseqm_b200.distributed import sp2_tile_ranges
    for ntile in (1, 3, 10, 1128, 4560):            # T = 1, 2, 4 (C60), 47 (config 5), 95
        for world in (1, 2, 3, 4, 8):
            for nsub in (1, 2, 4, 8):
                rng = sp2_tile_ranges(ntile, world, nsub)
                assert rng[0][0] == 0 and rng[-1][1] == ntile
                assert all(rng[k][1] == rng[k + 1][0] for k in range(world - 1))
                buf = (ctypes.c_int32 * 4)()
                _lib.lib().seqm_sp2_shard_plan(ntile, world, nsub, 0, buf)
                tpr, tps = buf[0], buf[1]
                assert tpr == tps * nsub and tpr * world >= ntile and (tpr - nsub) * world < ntile + nsub * world
                assert all(hi - lo <= tpr for lo, hi in rng)


def _plan_worker(rank, world, port, ntile, q):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from pyseqm_b200.distributed import sp2_tile_ranges
    lo, hi = sp2_tile_ranges(ntile, world)[rank]
    mine = torch.zeros(ntile, dtype=torch.int64)
    mine[lo:hi] = 1                                          # tiles this rank would multiply
    dist.all_reduce(mine)                                    # stands for the all-gather of the tile records
    q.put((rank, lo, hi, bool((mine == 1).all())))
    dist.destroy_process_group()


def test_sp2_tile_plan_world2_gloo():
    world, ntile = 2, 1128
    port = _free_port()
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    procs = [ctx.Process(target=_plan_worker, args=(r, world, port, ntile, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(r[3] for r in res) and res[0][2] == res[1][1]


class _FakeForce:
    """stands for pyseqm_b200.basics.Force on the CPU: per-molecule results that depend on the molecule only"""

    def __call__(self, const, xyz, sp):
        from pyseqm_b200.batching import packed_sizes
        e = (xyz * (sp > 0).unsqueeze(2)).sum(dim=(1, 2)) + packed_sizes(sp).to(xyz.dtype)
        out = [None] * 11
        out[0], out[2], out[3], out[10] = -xyz, e, 2.0 * e, torch.zeros(sp.shape[0], dtype=torch.bool)
        return tuple(out)


def _sweep_worker(rank, world, port, nmol, q):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    import numpy as np
    from pyseqm_b200 import synth
    from pyseqm_b200.batching import BucketedBatch, packed_sizes
    from pyseqm_b200.distributed import gather_molecules, shard_bounds
    sp, xyz = synth.mixed_organics_sweep(nmol, seed=0, base=24)
    species, coords = torch.as_tensor(sp), torch.as_tensor(xyz)
    lo, hi = shard_bounds(nmol, world)[rank]
    bb = BucketedBatch(species[lo:hi], max_chunk=8)
    f, e, hf, nc = bb.run(_FakeForce(), None, coords[lo:hi])
    e_all = gather_molecules(e, nmol)
    want = (coords * (species > 0).unsqueeze(2)).sum(dim=(1, 2)) + packed_sizes(species).to(coords.dtype)
    ok = bool(torch.allclose(e_all, want, rtol=0, atol=1e-12)) and bool(torch.equal(f, -coords[lo:hi] * (species[lo:hi] > 0).unsqueeze(2)))
    q.put((rank, lo, hi, ok))
    dist.destroy_process_group()


def test_bucketed_sweep_sharded_world2_gloo():
    """config 4's N > 1 host path: contiguous molecule shards, a size-bucketed sweep per rank, results scattered back
    into shard order and gathered into the global order (world size 2, gloo, a CPU stand-in for Force)"""
    world, nmol = 2, 61
    port = _free_port()
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    procs = [ctx.Process(target=_sweep_worker, args=(r, world, port, nmol, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=180) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(r[3] for r in res) and res[0][2] == res[1][1] and res[1][2] == nmol
