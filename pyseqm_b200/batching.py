"""Size-bucketed sweeps over large batches of mixed molecules (BASELINE.json configs[3]: 65,536 organics of 5-30 atoms).

The reference pads every molecule of a batch to the largest one (seqm/basics.py:43-118: species [B, S], zero padded)
and runs one SCF over the padded batch.  Molecules are independent (every reduction of the path is per molecule,
energy.py:92-93,109-114), so a sweep may be cut freely.  `BucketedBatch` sorts the molecules by packed orbital count
m = 4 nHeavy + nHydro, cuts them into buckets at the given edges (so that each bucket runs the SCF form and the
eigensolver instantiation that fits its size: m <= 32 takes the one-launch warp kernel at the bucket's own N, larger
ones the CTA Jacobi), trims each bucket's padding to its own largest molecule, bounds the bucket size by a memory
cap, runs `Force` per bucket and scatters the results back into the caller's molecule order.  The bucket species
tensors are kept, so the species cache of `_engine.Prepared` hits on every later sweep over the same batch.
"""
import torch

from . import _engine

DEFAULT_EDGES = (8, 12, 16, 20, 24, 28, 32, 40, 48, 56, 64, 80, 96, 118)


def packed_sizes(species):
    """m = 4 * (#atoms with Z > 1) + (#hydrogens) per molecule (reference pack.py:7-13)"""
    return 4 * (species > 1).sum(dim=1) + (species == 1).sum(dim=1)


class BucketedBatch:
    def __init__(self, species, edges=DEFAULT_EDGES, max_chunk=16384):
        m = packed_sizes(species)
        natoms = (species > 0).sum(dim=1)
        order = torch.argsort(m, stable=True)
        ms = m[order].cpu()
        self.nmol, self.molsize = species.shape
        self.buckets = []                                   # (index tensor into the batch, trimmed species tensor)
        lo = 0
        bounds = list(edges) + [1 << 30]
        for hi_m in bounds:
            hi = int(torch.searchsorted(ms, torch.tensor(hi_m), right=True))
            start = lo
            while start < hi:
                stop = min(start + max_chunk, hi)
                idx = order[start:stop]
                S = max(int(natoms[idx].max()), 1)
                self.buckets.append((idx, species[idx][:, :S].contiguous()))
                start = stop
            lo = hi
        if _engine._PREP_CACHE_SIZE < len(self.buckets) + 2:      # every bucket keeps its cached topology
            _engine._PREP_CACHE_SIZE = len(self.buckets) + 2

    def summary(self):
        return [(int(idx.numel()), int(sp.shape[1]), int(packed_sizes(sp).max())) for idx, sp in self.buckets]

    def run(self, force_module, const, coordinates, workers=0):
        """-> force [B,S,3], Etot [B], Hf [B], notconverged [B] in the caller's molecule order.
        workers > 0: the buckets run concurrently, one host thread, CUDA stream and `Force` module per bucket (the
        SCF of the larger molecules synchronises its stream once per iteration, which leaves the GPU idle between
        the launches of a small bucket; independent buckets fill those gaps).  The ctypes calls release the GIL."""
        dev = coordinates.device
        force = torch.zeros(self.nmol, self.molsize, 3, dtype=coordinates.dtype, device=dev)
        Etot = torch.zeros(self.nmol, dtype=coordinates.dtype, device=dev)
        Hf = torch.zeros_like(Etot)
        nc = torch.zeros(self.nmol, dtype=torch.bool, device=dev)
        coords = coordinates.detach()

        def one(k, fm):
            idx, sp = self.buckets[k]
            S = sp.shape[1]
            out = fm(const, coords[idx][:, :S].contiguous(), sp)
            return out[0], out[2].detach(), out[3].detach(), out[10]

        if workers <= 0 or len(self.buckets) < 2:
            results = [one(k, force_module) for k in range(len(self.buckets))]
        else:
            from concurrent.futures import ThreadPoolExecutor
            if not hasattr(self, '_workers'):
                self._workers = [(type(force_module)(force_module.seqm_parameters).to(dev), torch.cuda.Stream(device=dev))
                                 for _ in self.buckets]
            main = torch.cuda.current_stream(dev)

            def job(k):
                fm, stream = self._workers[k]
                with torch.cuda.device(dev), torch.cuda.stream(stream):
                    stream.wait_stream(main)
                    res = one(k, fm)
                stream.synchronize()
                return res

            # largest buckets first: they are the long poles
            order = sorted(range(len(self.buckets)), key=lambda k: -self.buckets[k][0].numel() * self.buckets[k][1].shape[1] ** 2)
            with ThreadPoolExecutor(max_workers=min(workers, len(order))) as pool:
                futs = {k: pool.submit(job, k) for k in order}
                results = [futs[k].result() for k in range(len(self.buckets))]
        for (idx, sp), (f, e, h, c) in zip(self.buckets, results):
            force[idx, :sp.shape[1]] = f
            Etot[idx] = e
            Hf[idx] = h
            nc[idx] = c
        return force, Etot, Hf, nc
