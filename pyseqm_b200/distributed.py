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


_GATHER_BUF = {}      # (shape of the padded shard, dtype, device, world) -> (send buffer, receive buffer)


class GatheredMolecules:
    """handle of an asynchronous gather_molecules(): .wait() -> the global-order tensor"""

    def __init__(self, work, recv, bounds, longest):
        self._work, self._recv, self._bounds, self._longest = work, recv, bounds, longest

    def wait(self):
        if self._work is not None:
            self._work.wait()
            self._work = None
        if all(hi - lo == self._longest for lo, hi in self._bounds):
            return self._recv.reshape((-1,) + tuple(self._recv.shape[2:]))
        return torch.cat([self._recv[r, :hi - lo] for r, (lo, hi) in enumerate(self._bounds)], dim=0)


def gather_molecules(local, nmol, world=None, async_op=False, group=None):
    """per-molecule results of every rank -> the global batch order, on every rank.  ONE all_gather_into_tensor into a
    receive buffer that is allocated once per shard shape (shards may differ in length by one molecule: the short ones
    are padded in the send buffer and trimmed on the way out).  async_op=True returns a GatheredMolecules handle at
    once - the collective runs on NCCL's stream while the caller enqueues the next step - and .wait() orders the
    current stream behind it when the result is consumed."""
    world = dist.get_world_size(group) if world is None else world
    bounds = shard_bounds(nmol, world)
    longest = max(hi - lo for lo, hi in bounds)
    if local.shape[0] > longest:
        raise ValueError("local shard is longer than shard_bounds allows")
    tail = tuple(local.shape[1:])
    key = ((longest,) + tail, local.dtype, str(local.device), world)
    bufs = _GATHER_BUF.get(key)
    if bufs is None:
        bufs = (torch.zeros((longest,) + tail, dtype=local.dtype, device=local.device),
                torch.empty((world, longest) + tail, dtype=local.dtype, device=local.device))
        _GATHER_BUF[key] = bufs
    send, recv = bufs
    send[:local.shape[0]].copy_(local.detach())
    work = dist.all_gather_into_tensor(recv.reshape((world * longest,) + tail), send, group=group, async_op=True)
    handle = GatheredMolecules(work, recv, bounds, longest)
    return handle if async_op else handle.wait()


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
