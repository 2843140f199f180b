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
