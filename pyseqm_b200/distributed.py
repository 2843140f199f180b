"""Molecule sharding for one-process-per-GPU runs.  The SCF hot path has no exchange step between molecules
(reference: every tensor of seqm/basics.py carries the molecule as its leading, independent batch index), so the
batch is cut into contiguous molecule ranges, one per rank, and the only collective is the gather of per-molecule
results at the end of a step."""
import torch
import torch.distributed as dist


def shard_bounds(nmol, world):
    """[lo, hi) per rank: contiguous, sizes differ by at most one molecule, earlier ranks get the longer shards"""
    base, extra = divmod(int(nmol), int(world))
    bounds, lo = [], 0
    for r in range(world):
        hi = lo + base + (1 if r < extra else 0)
        bounds.append((lo, hi))
        lo = hi
    return bounds


def shard_batch(species, coordinates, rank=None, world=None):
    """this rank's molecules (views, no copy) and its [lo, hi) range in the global batch"""
    rank = dist.get_rank() if rank is None else rank
    world = dist.get_world_size() if world is None else world
    lo, hi = shard_bounds(species.shape[0], world)[rank]
    return species[lo:hi], coordinates[lo:hi], (lo, hi)


def gather_molecules(local, nmol, world=None):
    """per-molecule results of every rank -> the global batch order, on every rank.  Shards may differ in length
    by one molecule: the short ones are padded for the all_gather and trimmed afterwards."""
    world = dist.get_world_size() if world is None else world
    bounds = shard_bounds(nmol, world)
    longest = max(hi - lo for lo, hi in bounds)
    pad = longest - local.shape[0]
    if pad < 0:
        raise ValueError("local shard is longer than shard_bounds allows")
    buf = local.detach().contiguous()
    if pad:
        buf = torch.cat([buf, buf.new_zeros((pad,) + tuple(buf.shape[1:]))], dim=0)
    parts = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(parts, buf)
    return torch.cat([p[:hi - lo] for p, (lo, hi) in zip(parts, bounds)], dim=0)


# ---- single large system (BASELINE.json configs[4]): SP2 tiles sharded over the ranks -------------------------
_sp2_comm = None


def enable_sp2_sharding(group=None):
    """Every rank holds the whole system (Parser, integrals, Fock build and gradient run replicated); the SP2
    purification - the dense X @ X of reference SP2.py:47 - is split by 64x64 tiles over the ranks of `group`,
    exchanged with ncclAllGather inside the library (include/seqm_b200.h, seqm_sp2_density_nccl).  Call once after
    init_process_group with this rank's CUDA device current; the 128-byte NCCL id travels through `group`."""
    global _sp2_comm
    from . import _ops
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    box = [_ops.nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
    _sp2_comm = _ops.nccl_comm_create(box[0], rank, world)
    _ops.sp2_sharding(_sp2_comm, world)
    return _sp2_comm


def disable_sp2_sharding():
    global _sp2_comm
    from . import _ops
    _ops.sp2_sharding(None, 0)
    if _sp2_comm is not None:
        _ops.nccl_comm_destroy(_sp2_comm)
        _sp2_comm = None


def sp2_tile_ranges(ntile, world, nsub=4):
    """[lo, hi) of the upper-block-triangle tiles each rank multiplies (the library's own plan, host only)"""
    import ctypes
    from . import _lib
    out = []
    for r in range(world):
        buf = (ctypes.c_int32 * 4)()
        _lib.lib().seqm_sp2_shard_plan(int(ntile), int(world), int(nsub), r, buf)
        out.append((buf[2], buf[3]))
    return out
