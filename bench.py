#!/usr/bin/env python
"""Benchmark of the batched SCF hot path.  BASELINE.json metric: "SCF molecules/sec & XL-BOMD steps/sec".

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload all|benzene|c60xl|water|organics]

Default (`--workload all`): ONE JSON line.  Top level = BASELINE.json configs[1], AM1 SCF + forces for 1024 perturbed
benzenes per GPU (12 atoms, 30 orbitals, eigh), metric SCF molecules/s; one step = one `Force` call (pair pass + SCF +
force adjoint) over the whole batch; `value` is timed with the coordinates resident in HBM, `e2e` includes, every step,
the host->device copy of the coordinates from pinned memory and the device->host copy of forces and energies.  N > 1:
one process per GPU (torchrun), every rank runs its own 1024-molecule shard (weak scaling), no collective in the loop,
one asynchronous gather of the energies per step.
"xlbomd" = configs[2]: XL-BOMD (k = 5) of 256 C60 per GPU, PM3, SP2: batch steps/s, with its own roofline / e2e /
cpu_baseline.  With N > 1 (or SEQM_BENCH_SUB=all) also "organics" = configs[3] (65,536 mixed organics, STRONG scaling by
molecule) and "water" = configs[4] ((H2O)500, SP2 tiles sharded over the GPUs, checked against the reference fixture).
`--impl reference`: the reference's own CPU path (oracle/_ref = the unmodified package, built by oracle/build_ref.sh;
else the oracle port) on the host cores, same configuration as the top-level workload.
"""
import argparse
import datetime
import json
import os
import subprocess
import sys
import threading
import time



# SEQM_BENCH_CPU=0 skips the CPU-baseline legs (kernel experiments on the GPU box; the driver's runs keep them)
CPU_LEG = os.environ.get('SEQM_BENCH_CPU', '1') != '0'


def host_cores():
    """host threads this process may really use: the scheduler affinity, capped by the cgroup CPU quota (a container
    that sees 24 CPUs but is limited to fewer runs a 24-thread OpenMP region 40x slower than one sized to its quota)"""
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    try:
        with open('/sys/fs/cgroup/cpu.max') as f:
            quota, period = f.read().split()[:2]
        if quota != 'max':
            n = min(n, max(1, int(float(quota) / float(period))))
    except Exception:
        try:
            with open('/sys/fs/cgroup/cpu/cpu.cfs_quota_us') as f:
                q = int(f.read())
            with open('/sys/fs/cgroup/cpu/cpu.cfs_period_us') as f:
                per = int(f.read())
            if q > 0:
                n = min(n, max(1, q // per))
        except Exception:
            pass
    return max(1, n)


# torchrun exports OMP_NUM_THREADS=1; the CPU legs (rank 0) must use the host cores the line reports
if os.environ.get('OMP_NUM_THREADS') == '1' and int(os.environ.get('RANK', '0')) == 0:
    os.environ['OMP_NUM_THREADS'] = str(host_cores())

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
T_START = time.time()


def log(msg):
    """progress on stderr (the JSON line is the only thing on stdout)"""
    if int(os.environ.get('RANK', '0')) == 0:
        sys.stderr.write("[bench %7.1fs] %s\n" % (time.time() - T_START, msg))
        sys.stderr.flush()
sys.path.insert(0, ROOT)
torch.set_default_dtype(torch.float64)

NMOL = 1024
WORKLOAD_BENZENE = ("AM1 SCF+forces, 1024 perturbed benzenes per GPU (12 atoms, 30 orbitals), eigh, converger [2], "
                    "scf_eps 1e-6 (BASELINE.json configs[1])")
SEQM_PAR = dict(method='AM1', scf_eps=1.0e-6, scf_converger=[2, 0.0], sp2=[False], elements=[0, 1, 6], learned=[],
                pair_outer_cutoff=1.0e10)


def peaks():
    try:
        with open(os.path.join(ROOT, 'MEASURED_PEAKS.json')) as f:
            return float(json.load(f)['hbm_gbs']), 'measured (MEASURED_PEAKS.json)'
    except Exception:
        return 6650.0, 'fallback (B200_PROFILING.md)'


class ClockSampler:
    """nvidia-smi clocks / throttle reasons while the timed region runs: ONE long-running nvidia-smi process
    (-lms 20, time-stamped rows) started before the warm-up (its start-up takes longer than a short timed region)
    and stopped after the timed region; only rows stamped inside [mark_begin, mark_end] are summarised."""

    def __init__(self, index):
        self.index, self.proc, self.rows, self.t0, self.t1 = index, None, [], None, None

    def start(self):
        q = 'timestamp,clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,' \
            'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,' \
            'clocks_event_reasons.sw_power_cap'
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.index), '--query-gpu=' + q,
                                          '--format=csv,noheader,nounits', '-lms', '20'], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def mark_begin(self):
        self.t0 = datetime.datetime.now()

    def mark_end(self):
        self.t1 = datetime.datetime.now()

    def stop(self):
        if self.proc is None:
            return
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:
            self.proc.kill()
            out = ''
        for line in out.strip().splitlines():
            t = [x.strip() for x in line.split(',')]
            if len(t) >= 7:
                try:
                    ts = datetime.datetime.strptime(t[0], '%Y/%m/%d %H:%M:%S.%f')
                    float(t[1])
                except ValueError:
                    continue
                self.rows.append((ts, t[1:]))

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        inside = [r for ts, r in self.rows if self.t0 is not None and self.t0 <= ts <= (self.t1 or ts)]
        window = "timed region"
        if not inside:      # region shorter than the sampling period: the rows under the same load just before it
            inside, window = [r for ts, r in self.rows], "warm-up + timed region"
        sm = sorted(float(r[0]) for r in inside)
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        reasons = [n for k, n in enumerate(names) if any(r[2 + k].lower().startswith('active') for r in inside)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(inside[0][1]), "reasons": reasons,
                "samples": len(inside), "window": window}


def load_reference():
    """the unmodified reference package from oracle/_ref (built by oracle/build_ref.sh), or None"""
    root = os.path.join(ROOT, 'oracle', '_ref')
    if not os.path.exists(os.path.join(root, 'seqm', 'basics.py')):
        return None
    import warnings
    warnings.filterwarnings('ignore', category=SyntaxWarning)
    if root not in sys.path:
        sys.path.insert(0, root)
    import seqm
    return seqm if os.path.abspath(seqm.__file__).startswith(root) else None


class CpuPath:
    """The reference's CPU implementation of the path on this box's host cores: the unmodified reference (oracle/_ref,
    kind "reference") when the recipe has installed it, else the oracle port (kind "port")."""

    def __init__(self):
        import warnings
        warnings.filterwarnings('ignore')
        self.seqm = load_reference()
        self.kind = 'reference' if self.seqm is not None else 'port'
        self.cores = host_cores()
        torch.set_num_threads(self.cores)
        if self.seqm is None:
            sys.path.insert(0, os.path.join(ROOT, 'oracle'))
            import seqm_oracle
            self.orc = seqm_oracle

    def _par(self, method, sp2, elements):
        p = dict(method=method, scf_eps=1.0e-6, scf_converger=[2, 0.0], sp2=list(sp2), elements=list(elements), learned=[],
                 pair_outer_cutoff=1.0e10)
        p['parameter_file_dir'] = os.path.join(os.path.dirname(self.seqm.__file__), 'params') + '/'
        return p

    def force_seconds(self, sp, xyz, method, sp2=(False,), elements=(0, 1, 6, 7, 8)):
        """wall time of one Force call (SCF + forces) over the batch"""
        species, coords = torch.as_tensor(sp), torch.as_tensor(xyz).clone()
        t0 = time.time()
        if self.seqm is not None:
            from seqm.basics import Force
            from seqm.seqm_functions.constants import Constants
            Force(self._par(method, sp2, elements))(Constants(), coords, species)
        else:
            self.orc.run(species, coords, method=method, scf_eps=1e-6, converger=(2, 0.0), sp2=tuple(sp2))
        return time.time() - t0

    def xlbomd_seconds_per_step(self, sp, xyz, steps):
        """XL-BOMD (k = 5, PM3, SP2 1e-5) steady-state seconds per step over the batch; None for the port (the oracle has
        no MD driver)"""
        if self.seqm is None:
            return None
        from seqm.XLBOMD import XL_BOMD
        from seqm.seqm_functions.constants import Constants
        par = self._par('PM3', [True, 1.0e-5], [0, 6])
        out = {'molid': [0], 'thermo': 10 ** 9, 'dump': 10 ** 9, 'prefix': '/tmp/seqm_ref_bench'}
        const = Constants()
        species, coords = torch.as_tensor(sp), torch.as_tensor(xyz).clone()
        md = XL_BOMD(par, timestep=1.0, k=5, output=out)
        torch.manual_seed(0)
        vel = md.initialize_velocity(const, coords, species, Temp=300.0)
        import contextlib, io
        with contextlib.redirect_stdout(io.StringIO()):
            md.run(const, 1, coords, vel, species)              # initialisation (full SCF) + one step: untimed
            t0 = time.time()
            md.run(const, steps + 1, coords, vel, species)      # SCF again + steps + 1 steps
            t1 = time.time()
            md.run(const, 1, coords, vel, species)
            t2 = time.time()
        return ((t1 - t0) - (t2 - t1)) / steps                  # the initialisation cancels


def run_reference(args, rank, world):
    """`--impl reference`: the reference's own CPU path on the host cores, SAME configuration as the product arm (1024
    perturbed benzenes per step, AM1, eigh, converger [2], scf_eps 1e-6), every step one Force call over all of them;
    plus the XL-BOMD half of the metric on a bounded sample (2 C60)."""
    if rank != 0:
        return
    from pyseqm_b200 import synth
    cpu = CpuPath()
    sp, xyz = synth.benzene(NMOL, seed=0)
    # every step is the full 1024-molecule batch (~3 s on 16 cores): at most 20 timed steps and 3 warm-up calls, so that
    # the arm ends within a few minutes whatever K and W are (the line reports the steps it really timed)
    nwarm, nstep = min(args.warmup, 3), min(args.steps, 20)
    times = []
    for it in range(nwarm + nstep):
        dt = cpu.force_seconds(sp, xyz, 'AM1', elements=(0, 1, 6))
        if it >= nwarm:
            times.append(dt)
    ms = 1e3 * sum(times) / len(times)
    val = NMOL / (ms * 1e-3)
    sample = "all %d perturbed benzenes of the product arm's rank 0 per step (same generator, seed 0)" % NMOL
    spc, xyzc = synth.c60(2, seed=0)
    xl = cpu.xlbomd_seconds_per_step(spc, xyzc, 3)
    line = {
        "impl": "reference", "metric": "SCF molecules/sec", "value": val, "unit": "molecules/s", "n_gpus": args.gpus,
        "steps": nstep, "warmup": nwarm, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD_BENZENE, "molecules_per_step": NMOL, "sample": sample},
        "cpu_baseline": {"value": val, "unit": "molecules/s", "cores": cpu.cores, "kind": cpu.kind, "sample": sample},
        "e2e": {"value": val, "unit": "molecules/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    if xl:
        line["xlbomd"] = {"metric": "XL-BOMD steps/sec", "value": 1.0 / xl * (2.0 / 256.0), "unit": "batch steps/s (256 C60 per batch)",
                          "sample": "2 C60, 3 steady-state steps, scaled to the 256-molecule batch of the product arm "
                                    "(cost is linear in the number of molecules)", "kind": cpu.kind, "cores": cpu.cores,
                          "seconds_per_step_2_c60": xl}
    print(json.dumps(line))


def ncu_traffic(key):
    """DRAM bytes per launch of a kernel from the committed ncu --set full capture of this round (profiles/
    ncu_traffic.json, written by tools/ncu_summary.py from the .ncu-rep); None if there is no capture"""
    try:
        with open(os.path.join(ROOT, 'profiles', 'ncu_traffic.json')) as f:
            e = json.load(f).get(key)
        return (int(e['dram_bytes']), e['source']) if e else (None, None)
    except Exception:
        return None, None


def profile_report(L):
    import ctypes
    buf = ctypes.create_string_buffer(1 << 16)
    L.seqm_profile_report(buf, len(buf))
    L.seqm_profile_enable(0)
    prof = {}
    for line in buf.value.decode().splitlines():
        nm, ms, cnt = line.rsplit(' ', 2)
        prof[nm] = {"ms": float(ms), "launch_groups": int(cnt)}
    tot = sum(v["ms"] for v in prof.values()) or 1.0
    for v in prof.values():
        v["share"] = v["ms"] / tot
    return prof, tot


def dgemm_peak(dev):
    """FP64 GEMM rate of this box, measured here (cuBLAS DGEMM through torch.matmul, 4096^3, best of 5): the tensor
    roofline's denominator (MEASURED_PEAKS.json holds no FP64 number)"""
    A_ = torch.randn(4096, 4096, device=dev)
    B_ = torch.randn(4096, 4096, device=dev)
    torch.matmul(A_, B_)
    best = 1e30
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        torch.matmul(A_, B_)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return 2.0 * 4096 ** 3 / (best * 1e-3) / 1e12


class Ctx:
    """one process of the run: rank, device, library, distributed helpers (the process group is created once in main)"""

    def __init__(self, args, rank, world, local_rank):
        self.args, self.rank, self.world = args, rank, world
        self.dev = torch.device('cuda', local_rank)
        self.local_rank = local_rank

    def barrier(self):
        if self.world > 1:
            import torch.distributed as dist
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(self, vals):
        t = torch.tensor([float(v) for v in vals], device=self.dev)
        if self.world > 1:
            import torch.distributed as dist
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return [float(x) for x in t]


def algorithmic_bytes(topo):
    """bytes one launch group must move at least once, per kernel"""
    nHH, nXH, nXX = topo.ncls
    A, B, S, m = topo.natoms, topo.nmol, topo.molsize, topo.mmax
    blk = {0: 8, 1: 32, 2: 128}              # bytes of one P/M/F block of the class (1, 4, 16 doubles)
    wb = {0: 8, 1: 80, 2: 800}               # integrals of one pair
    pv = {0: 16, 1: 88, 2: 160}              # Coulomb vectors written per pair
    n = {0: nHH, 1: nXH, 2: nXX}
    fock_pair = sum(n[c] * (wb[c] + 2 * blk[c] + 2 * blk[c] + pv[c] + 16) for c in range(3))
    fock_diag = A * (80 + 80 + 2 * 80) + sum(n[c] * pv[c] for c in range(3))
    pair_fwd = sum(n[c] * (32 + 16 + wb[c] + 128 + 8 + pv[c]) for c in range(3)) + A * 224
    pair_grad = sum(n[c] * (32 + 16 + 3 * blk[c] + 32) for c in range(3)) + A * 224
    eigh = B * 2 * m * m * 8
    # one-launch SCF (K3s): integrals + core Hamiltonian + initial density in, density + Fock matrix out, once
    scf_small = sum(n[c] * wb[c] for c in range(3)) + B * 4 * m * m * 8 + A * 40
    return dict(scf_small=scf_small, fock_pair=fock_pair, fock_diag=fock_diag, fock=fock_pair + fock_diag, pair_fwd=pair_fwd,
                pair_grad=pair_grad, eigh_density=eigh)


def time_op(fn, flush, reps=5):
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    fn()
    for a, b in ev:
        flush()
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    return sum(a.elapsed_time(b) for a, b in ev) / reps


def run_gpu(ctx):
    """BASELINE.json configs[1]: the headline workload.  -> result dict on rank 0, None elsewhere"""
    from pyseqm_b200 import _lib, _ops, synth
    from pyseqm_b200.basics import Force, GraphedForce
    from pyseqm_b200.seqm_functions.constants import Constants
    args, rank, world, dev = ctx.args, ctx.rank, ctx.world, ctx.dev
    use_graph = os.environ.get('SEQM_BENCH_GRAPH', '1') != '0'
    if world > 1:
        import torch.distributed as dist
        from pyseqm_b200.distributed import gather_molecules
    L = _lib.lib()
    sp, xyz = synth.benzene(NMOL, seed=rank)
    species = torch.as_tensor(sp).to(dev)
    coords_host = torch.as_tensor(xyz).pin_memory()
    coords = coords_host.to(dev)
    const = Constants().to(dev)
    # defer_scf_warning: the SCF status of step k is read when step k+1 starts, so the host enqueues the next step
    # while the GPU still runs this one (pyseqm_b200/basics.py:Force)
    force_mod = Force(dict(SEQM_PAR, defer_scf_warning=True)).to(dev)
    flush_buf = torch.empty(192 * 1024 * 1024, dtype=torch.uint8, device=dev)     # > 126 MB of L2

    def flush():
        flush_buf.fill_(1)

    pending = []

    graphed = GraphedForce(dict(SEQM_PAR)).to(dev) if use_graph else None

    def step(x, eager=False):
        # one Force call over the batch: a CUDA-graph replay of the call (pyseqm_b200.basics.GraphedForce) unless eager
        out = force_mod(const, x, species) if (eager or graphed is None) else graphed(const, x, species)
        if world > 1:
            # energies of every shard in global molecule order: one preallocated all_gather_into_tensor, asynchronous on
            # NCCL's stream; the result of step k is consumed (waited for) while step k+1 is already enqueued
            pending.append(gather_molecules(out[2], NMOL * world, world, async_op=True))
            if len(pending) > 1:
                pending.pop(0).wait()
        return out

    def drain():
        while pending:
            pending.pop(0).wait()
        force_mod.flush_status()

    barrier = ctx.barrier
    local_rank = ctx.local_rank
    # cold call: first call with this species tensor (Parser, parameter gather, topology, per-atom records, initial
    # density: everything _engine.Prepared caches afterwards), after the CUDA context / module load of a dry call
    Force(dict(SEQM_PAR)).to(dev)(const, coords[:8].clone(), species[:8].clone())
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = step(coords.clone(), eager=True)
    torch.cuda.synchronize()
    cold_ms = (time.perf_counter() - t0) * 1e3
    # library launches of one call (a graph replay re-issues exactly the launches of the captured call; the library's
    # counter only sees the capture)
    force_mod.flush_status()
    lc0 = L.seqm_launch_count()
    step(coords.clone(), eager=True)
    drain()
    launches_per_call = int(L.seqm_launch_count() - lc0)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    for _ in range(args.warmup):
        out = step(coords.clone())
    drain()
    nconv_fail = int(out[10].sum())
    # ---- resident-input timing ----
    barrier()
    sampler.mark_begin()
    l0 = L.seqm_launch_count()
    evs = []
    for _ in range(args.steps):
        x = coords.clone()
        flush()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        step(x)
        b.record()
        evs.append((a, b))
    drain()
    barrier()
    launches = int(L.seqm_launch_count() - l0) if graphed is None else launches_per_call * args.steps
    total_ms = sum(a.elapsed_time(b) for a, b in evs)
    log("benzene: per-step ms " + " ".join("%.3f" % a.elapsed_time(b) for a, b in evs))
    # ---- end to end: pinned host coordinates in, forces + energies out, every step ----
    f_host = torch.empty(NMOL, 12, 3).pin_memory()
    e_host = torch.empty(NMOL).pin_memory()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        x = coords_host.to(dev, non_blocking=True)
        o = step(x)
        f_host.copy_(o[0], non_blocking=True)
        e_host.copy_(o[2].detach(), non_blocking=True)
        torch.cuda.synchronize()
    drain()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    sampler.mark_end()
    if rank == 0:
        sampler.stop()
    total_ms, e2e_ms = ctx.max_over_ranks([total_ms, e2e_ms])
    if rank != 0:
        return None
    # ---- per-kernel profile of one step, and the kernels timed alone on the converged state ----
    L.seqm_profile_enable(1)
    force_mod(const, coords.clone(), species)        # no collective here: the other ranks have finished
    force_mod.flush_status()
    prof, tot_prof = profile_report(L)
    st = force_mod.energy.last_state
    topo = st['topo']
    ab = algorithmic_bytes(topo)
    hbm, peak_src = peaks()
    xij = torch.randn(topo.npairs, 3, device=dev)
    xij = xij / xij.norm(dim=1, keepdim=True)
    rij = torch.rand(topo.npairs, device=dev) * 6 + 2
    par = force_mod.energy.packpar(topo.Z.long(), learned_params=dict())
    ops = {
        'fock': lambda: _ops.fock(topo, st['w'], st['P'], st['M'], st['par5'], pv=st['pv'], F=st['F']),
        'eigh_density': lambda: _ops.eigh_density(st['F'], topo.nheavy, topo.nhydro, topo.nocc, topo.molsize, topo.mmax),
        'pair_fwd': lambda: _ops.hcore(topo, st['atompar'], xij, rij, par['U_ss'], par['U_pp'], 'AM1'),
        'pair_grad': lambda: _ops.pair_grad(topo, st['atompar'], xij, rij, 'AM1', st['P'], None, None, None),
    }
    kernels = {}
    for nm, fn in ops.items():
        ms = time_op(fn, flush)
        kernels[nm] = {"ms_per_launch": ms, "algorithmic_bytes": ab[nm], "achieved_gbs": ab[nm] / (ms * 1e-3) / 1e9,
                       "frac_of_hbm": ab[nm] / (ms * 1e-3) / 1e9 / hbm}
    if 'scf_small' in prof:      # a single launch per step: its duration inside the step is the measurement
        ms = prof['scf_small']['ms'] / max(prof['scf_small']['launch_groups'], 1)
        kernels['scf_small'] = {"ms_per_launch": ms, "algorithmic_bytes": ab['scf_small'],
                                "achieved_gbs": ab['scf_small'] / (ms * 1e-3) / 1e9,
                                "frac_of_hbm": ab['scf_small'] / (ms * 1e-3) / 1e9 / hbm,
                                "note": "latency/FP64-issue bound: one warp per molecule, 6.9 warps per SM (DESIGN.md)"}
    group = {'fock': ('fock_pair', 'fock_diag'), 'eigh_density': ('eigh_density',), 'pair_fwd': ('pair_fwd',),
             'pair_grad': ('pair_grad',), 'scf_small': ('scf_small',)}
    share = {k: sum(prof.get(g, {"share": 0.0})["share"] for g in gs) for k, gs in group.items()}
    top = max(share, key=share.get)
    roof = {"bound": "hbm", "kernel": top, "share_of_step": share[top], "achieved": kernels[top]["achieved_gbs"],
            "peak": hbm, "peak_source": peak_src, "unit": "GB/s", "frac": kernels[top]["frac_of_hbm"], "traffic": None}
    if top == 'scf_small':
        # ncu --set full of this kernel on this workload (profiles/r01p_scf_small_ncu.txt): DRAM bytes per launch and
        # what actually bounds it - neither roofline: 6.9 warps per SM, FP64 pipe 19 % busy, issue slots 26 % busy
        m_, B_ = int(topo.mmax), int(topo.nmol)
        iters = int(st['iterations']) + 1
        flops = B_ * iters * (9.0 * m_ ** 3 + 2.0 * m_ * m_ * int(topo.nocc.max()))      # 9 m^3 + 2 m^2 nocc per eigensolve
        traffic, tsrc = ncu_traffic('scf_small_benzene1024')
        roof.update({"traffic": traffic, "traffic_source": tsrc,
                     "limiter": "latency of one warp per molecule (6.9 molecules per SM): the dependent chain of a ring-Jacobi "
                                "round (dot product -> reduction -> tangent -> FMAs -> shuffles, ~600-800 cycles) and "
                                "the L2 latency of the Fock / Pulay phases; ncu r02g/h: issue slots 34 %, FP64 pipe 22 %, "
                                "stalls wait / long_scoreboard - DESIGN.md section 4",
                     "algorithmic_tflops": flops / (kernels[top]["ms_per_launch"] * 1e-3) / 1e12,
                     "algorithmic_flops_note": "(9 m^3 + 2 m^2 nocc) per eigensolve x (SCF iterations + 1) x molecules"})
    # ---- CPU baseline (rank 0, N = 1 only): the reference's CPU path (oracle/_ref, else the oracle port) on this batch ----
    cpu_base = None
    if world == 1 and CPU_LEG:
        log("benzene: cpu baseline")
        cpu = CpuPath()
        ns, reps = NMOL, 2
        cpu.force_seconds(sp[:64], xyz[:64], 'AM1', elements=(0, 1, 6))
        cpu_s = sum(cpu.force_seconds(sp, xyz, 'AM1', elements=(0, 1, 6)) for _ in range(reps)) / reps
        cpu_base = {"value": ns / cpu_s, "unit": "molecules/s", "cores": cpu.cores, "kind": cpu.kind,
                    "sample": "the %d benzenes of rank 0, %d Force calls" % (ns, reps)}
    ms_step = total_ms / args.steps
    nm_total = NMOL * world
    return ({
        "metric": "SCF molecules/sec", "value": nm_total / (ms_step * 1e-3), "unit": "molecules/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD_BENZENE,
                   "molecules_per_gpu": NMOL, "l2": "192 MiB buffer written between timed steps",
                   "scf_not_converged": nconv_fail, "scf_iterations": int(st['iterations']),
                   "api": ("pyseqm_b200.basics.GraphedForce: one Force call (pair pass, one-launch SCF, adjoint, energies) "
                           "replayed as one CUDA graph per step; gpu_launches = kernels of the captured call x steps"
                           if graphed is not None else
                           "pyseqm_b200.basics.Force with defer_scf_warning=True (SCF status of a call read at the next "
                           "call)") + "; N > 1: asynchronous gather of the energies, consumed one step later"},
        "cold_call_ms": cold_ms,
        "cold_call_note": "first Force call with a new species tensor (Parser, parameter gather, topology, per-atom "
                          "records, initial guess built and cached), CUDA context already up; steady state is ms_per_step",
        "e2e": {"value": nm_total / (e2e_ms / args.steps * 1e-3), "unit": "molecules/s",
                "h2d_bytes_per_step": NMOL * 12 * 3 * 8, "d2h_bytes_per_step": NMOL * 12 * 3 * 8 + NMOL * 8},
        "gpu_launches": launches, "clocks": sampler.summary(), "roofline": roof, "kernels": kernels,
        "profile_one_step": prof,
        "cpu_baseline": cpu_base})


def run_c60xl(ctx):
    """Secondary workload (BASELINE.json configs[2]): XL-BOMD (k=5) of 256 C60 per GPU, PM3, SP2 density
    (eps 1e-5); one step = one `XL_BOMD.one_step` of the whole batch (pair pass, Fock build with the dynamical
    density, one SP2 purification, shadow energy, force adjoint, Verlet).  metric: batch MD steps per second."""
    from pyseqm_b200 import _lib, synth
    from pyseqm_b200.XLBOMD import XL_BOMD
    from pyseqm_b200.seqm_functions.constants import Constants
    nmol = int(os.environ.get('SEQM_BENCH_C60', '256'))
    args, rank, world, dev, local_rank = ctx.args, ctx.rank, ctx.world, ctx.dev, ctx.local_rank
    L = _lib.lib()
    par = dict(method='PM3', scf_eps=1.0e-6, scf_converger=[2, 0.0], sp2=[True, 1.0e-5], elements=[0, 6], learned=[],
               pair_outer_cutoff=1.0e10, xl_cuda_graph=os.environ.get('SEQM_XL_GRAPH', '1') != '0')
    out = {'molid': [0], 'thermo': 10 ** 9, 'dump': 10 ** 9, 'prefix': '/tmp/seqm_b200_bench'}
    sp, xyz = synth.c60(nmol, seed=rank)
    species, coords = torch.as_tensor(sp).to(dev), torch.as_tensor(xyz).to(dev)
    const = Constants().to(dev)
    md = XL_BOMD(par, timestep=1.0, k=5, output=out).to(dev)
    torch.manual_seed(rank)
    vel = md.initialize_velocity(const, coords, species, Temp=300.0)
    mass = md._masses(const, species)
    acc, D = md.initialize(const, mass, coords, species)
    with torch.no_grad():
        Pt = D.unsqueeze(0).expand((md.m,) + D.shape).clone()
        P = D.clone()
    state = [coords, vel, acc, D, P, Pt]
    stepno = [0]

    def step():
        c, v, a, d, p_, pt = state
        c, v, a, d, p_, pt, Hf, force = md.one_step(const, stepno[0], mass, c, v, species, a, d, p_, pt)
        state[:] = [c, v, a, d, p_, pt]
        stepno[0] += 1
        return c, Hf

    barrier = ctx.barrier
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    for _ in range(max(args.warmup, 3)):      # step 1 calibrates the SP2 step count, step 2 captures the graph
        step()
    barrier()
    sampler.mark_begin()
    l0 = L.seqm_launch_count()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(args.steps):
        step()
    b.record()
    barrier()
    xs = getattr(md, '_xlstep', None)
    graph = xs is not None and xs.graph is not None
    # a replayed graph launches the kernels counted once at capture time
    launches = int(xs.launches_per_step * args.steps) if graph else int(L.seqm_launch_count() - l0)
    total_ms = a.elapsed_time(b)
    c_host = torch.empty(nmol, 60, 3).pin_memory()
    e_host = torch.empty(nmol).pin_memory()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        c, Hf = step()
        c_host.copy_(c.detach(), non_blocking=True)
        e_host.copy_(Hf.detach(), non_blocking=True)
        torch.cuda.synchronize()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    sampler.mark_end()
    if rank == 0:
        sampler.stop()
    total_ms, e2e_ms = ctx.max_over_ranks([total_ms, e2e_ms])
    if rank != 0:
        return None
    # per-kernel profile: one step with the launches enqueued one by one (a replayed graph has no event marks)
    if xs is not None:
        xs.want_graph, keep = False, xs.graph
    L.seqm_profile_enable(1)
    step()
    prof, tot_prof = profile_report(L)
    if xs is not None:
        xs.want_graph, xs.graph = graph, keep
    fp64_peak = dgemm_peak(dev)
    m = 240
    sp = prof.get('sp2_step', {"ms": 0.0, "launch_groups": 1, "share": 0.0})
    nstep_launch = xs.sp2_steps if xs is not None else 2 * max(sp["launch_groups"], 1)
    sp2_needed = nstep_launch - 4 if xs is not None else nstep_launch      # XLStep launches the calibrated count + 4 no-ops
    flops = 2.0 * m ** 3 * nmol                              # 2 m^3 per multiplication and molecule
    per_launch = sp["ms"] / max(sp2_needed, 1)               # the no-op launches cost microseconds
    T_ = (m + 63) // 64
    exec_frac = (T_ + 1) / (2.0 * T_)
    roof = {"bound": "tensor", "kernel": "sp2_step_kernel (X^2 on the upper block triangle, DMMA m8n8k4 f64, fused trace "
                                         "rule / update / transpose epilogue)", "share_of_step": sp["share"],
            "achieved": flops / (per_launch * 1e-3) / 1e12 if per_launch else None, "peak": fp64_peak,
            "peak_source": "cuBLAS DGEMM 4096^3 measured in this run (MEASURED_PEAKS.json holds no FP64 number)",
            "unit": "TFLOP/s", "frac": flops / (per_launch * 1e-3) / 1e12 / fp64_peak if per_launch else None,
            "executed_tflops": exec_frac * flops / (per_launch * 1e-3) / 1e12 if per_launch else None,
            "executed_frac": exec_frac * flops / (per_launch * 1e-3) / 1e12 / fp64_peak if per_launch else None,
            "traffic": ncu_traffic('sp2_step_c60x256')[0], "traffic_source": ncu_traffic('sp2_step_c60x256')[1],
            "note": "`achieved` / `frac` use the reference's count, 2 m^3 per product (SURVEY 8d); the kernel multiplies "
                    "only %d of %d tiles (symmetry): `executed_*` = the DMMA work really done over the same time, the "
                    "number to hold against the DGEMM rate" % (T_ * (T_ + 1) // 2, T_ * T_)}
    # HBM fraction of the integral / Fock kernels at this size (one launch each inside the profiled step)
    hbm, peak_src = peaks()
    npair = nmol * 60 * 59 // 2
    kb = {'pair_fwd': npair * (48 + 800 + 128 + 8 + 160) + nmol * 60 * 224,
          'fock_pair': npair * (800 + 4 * 128 + 160 + 16),
          'pair_grad': npair * (48 + 3 * 128 + 32) + nmol * 60 * 224}
    kernels = {}
    for nm_, nbytes in kb.items():
        if nm_ in prof and prof[nm_]["ms"] > 0:
            ms_ = prof[nm_]["ms"] / max(prof[nm_]["launch_groups"], 1)
            kernels[nm_] = {"ms_per_launch": ms_, "algorithmic_bytes": nbytes, "achieved_gbs": nbytes / (ms_ * 1e-3) / 1e9,
                            "frac_of_hbm": nbytes / (ms_ * 1e-3) / 1e9 / hbm}
    ms_step = total_ms / args.steps
    cpu_base, xl_s = None, None
    if world == 1 and CPU_LEG:      # CPU baselines are an N = 1 item
        log("xlbomd: cpu baseline")
        cpu = CpuPath()
        spc, xyzc = synth.c60(2, seed=0)
        xl_s = cpu.xlbomd_seconds_per_step(spc, xyzc, 3)
    if xl_s:
        cpu_base = {"value": 1.0 / xl_s * (2.0 / nmol), "unit": "batch steps/s", "cores": cpu.cores, "kind": cpu.kind,
                    "sample": "XL_BOMD of the reference on 2 C60, 3 steady-state steps (initialisation subtracted), scaled "
                              "to the %d-molecule batch (cost linear in the number of molecules)" % nmol}
    return ({
        "metric": "XL-BOMD steps/sec", "value": world / (ms_step * 1e-3), "unit": "batch steps/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "XL-BOMD k=5, %d C60 per GPU (60 atoms, 240 orbitals), PM3, SP2 eps 1e-5 "
                               "(BASELINE.json configs[2])" % nmol, "molecules_per_gpu": nmol,
                   "l2": "working set per step (P, F, D, w, ring: > 1 GB) exceeds L2",
                   "api": "pyseqm_b200.XLBOMD.XL_BOMD.one_step, fused step (%s)" % ("one CUDA graph replay per step" if graph else "launch by launch"),
                   "sp2_steps_launched": nstep_launch},
        "e2e": {"value": world / (e2e_ms / args.steps * 1e-3), "unit": "batch steps/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": nmol * 60 * 3 * 8 + nmol * 8,
                "note": "MD state lives on the device (as in the reference); coordinates and Hf go to pinned host "
                        "memory every step"},
        "gpu_launches": launches, "clocks": sampler.summary(), "roofline": roof, "kernels": kernels,
        "profile_one_step": prof, "cpu_baseline": cpu_base})


def run_water(ctx):
    """Third workload (BASELINE.json configs[4]): PM3 SCF + forces of ONE (H2O)_500 cluster (1500 atoms, 3000 packed
    orbitals), SP2 density (eps 1e-5).  One step = one `Force` call.  N > 1: STRONG scaling - every rank holds the
    whole system (pair pass, Fock build and gradient replicated, < 10 % of the step), the SP2 multiplications are
    tile-sharded over the ranks with ncclAllGather (pyseqm_b200.distributed.enable_sp2_sharding)."""
    from pyseqm_b200 import _lib, synth
    from pyseqm_b200.basics import Force
    from pyseqm_b200.seqm_functions.constants import Constants
    nwater = int(os.environ.get('SEQM_BENCH_WATER', '500'))
    args, rank, world, dev, local_rank = ctx.args, ctx.rank, ctx.world, ctx.dev, ctx.local_rank
    if world > 1:
        import torch.distributed as dist
        from pyseqm_b200.distributed import disable_sp2_sharding, enable_sp2_sharding
        enable_sp2_sharding()
    L = _lib.lib()
    if os.environ.get('SEQM_SP2_P2P'):
        L.seqm_sp2_nccl_p2p(int(os.environ['SEQM_SP2_P2P']))
    if os.environ.get('SEQM_SP2_NSUB'):
        L.seqm_sp2_nccl_subchunks(int(os.environ['SEQM_SP2_NSUB']))
    par = dict(method='PM3', scf_eps=1.0e-6, scf_converger=[2, 0.0], sp2=[True, 1.0e-5], elements=[0, 1, 8], learned=[],
               pair_outer_cutoff=1.0e10)
    sp, xyz = synth.water_cluster(nwater, seed=0)
    species = torch.as_tensor(sp).to(dev)
    coords_host = torch.as_tensor(xyz).pin_memory()
    coords = coords_host.to(dev)
    const = Constants().to(dev)
    fm = Force(par).to(dev)
    barrier = ctx.barrier
    nsteps = max(1, min(args.steps, int(os.environ.get('SEQM_BENCH_WATER_STEPS', '3'))))
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    for _ in range(min(args.warmup, 2)):
        out = fm(const, coords.clone(), species)
    barrier()
    sampler.mark_begin()
    l0 = L.seqm_launch_count()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(nsteps):
        out = fm(const, coords.clone(), species)
    b.record()
    barrier()
    launches = int(L.seqm_launch_count() - l0)
    total_ms = a.elapsed_time(b)
    f_host = torch.empty(1, 3 * nwater, 3).pin_memory()
    e_host = torch.empty(1).pin_memory()
    barrier()
    t0 = time.perf_counter()
    for _ in range(nsteps):
        o = fm(const, coords_host.to(dev, non_blocking=True), species)
        f_host.copy_(o[0], non_blocking=True)
        e_host.copy_(o[2].detach(), non_blocking=True)
        torch.cuda.synchronize()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    sampler.mark_end()
    if rank == 0:
        sampler.stop()
    total_ms, e2e_ms = ctx.max_over_ranks([total_ms, e2e_ms])
    # per-kernel profile of one more call: collective inside, so every rank runs it; rank 0 reports
    L.seqm_profile_enable(1 if rank == 0 else 0)
    c0 = L.seqm_sp2_step_count()
    fm(const, coords.clone(), species)
    torch.cuda.synchronize()
    sp2_steps_profiled = L.seqm_sp2_step_count() - c0
    st = fm.energy.last_state
    if world > 1:
        dist.barrier()
        disable_sp2_sharding()
    # parity of the sharded run: every rank holds the whole system, so rank 0's energy must equal the single-GPU
    # reference fixture of the same geometry (tests/golden_big/water500_PM3_sp2_c2.npz: seed 0, 500 waters)
    parity = None
    fx = os.path.join(ROOT, 'tests', 'golden_big', 'water500_PM3_sp2_c2.npz')
    if nwater == 500 and os.path.exists(fx) and rank == 0:
        g = np.load(fx)
        parity = {"max_abs_dEtot_eV_per_water": float(np.abs(out[2].detach().cpu().numpy() - g['Etot']).max() / nwater),
                  "max_abs_dforce_eV_per_A": float(np.abs(out[0].cpu().numpy() - g['force']).max()),
                  "against": "tests/golden_big/water500_PM3_sp2_c2.npz (unmodified reference, CPU)"}
    if rank != 0:
        return None
    prof, tot_prof = profile_report(L)
    fp64_peak = dgemm_peak(dev)
    m = int(st['topo'].mmax)
    spp = prof.get('sp2_step', {"ms": 0.0, "launch_groups": 1, "share": 0.0})
    nstep = max(int(sp2_steps_profiled), 1)                  # purification steps of the profiled call (library counter)
    flops = 2.0 * m ** 3                                     # 2 m^3 per multiplication
    per_step = spp["ms"] / nstep
    T_ = (m + 63) // 64
    ach = flops / (per_step * 1e-3) / 1e12 if per_step else None
    roof = {"bound": "tensor", "kernel": "sp2_step_kernel%s" % ("<sharded> + ncclAllGather + sp2_unpack_kernel" if world > 1 else ""),
            "share_of_step": spp["share"], "achieved": ach, "peak": fp64_peak * world,
            "peak_source": "cuBLAS DGEMM 4096^3 measured in this run on rank 0, times the number of GPUs",
            "unit": "TFLOP/s", "frac": ach / (fp64_peak * world) if ach else None,
            "executed_frac": (ach * (T_ + 1) / (2.0 * T_)) / (fp64_peak * world) if ach else None,
            "traffic": ncu_traffic('sp2_step_water500')[0] if world == 1 else None,
            "traffic_source": ncu_traffic('sp2_step_water500')[1] if world == 1 else None,
            "note": "2 m^3 flops per purification step over the time of one step (multiplication, all-gather and "
                    "unpack) on all GPUs; %d of %d tiles executed (symmetry)" % (T_ * (T_ + 1) // 2, T_ * T_)}
    ms_step = total_ms / nsteps
    return ({
        "metric": "SCF systems/sec (one 1500-atom system, SCF + forces)", "value": 1.0 / (ms_step * 1e-3), "unit": "Force calls/s",
        "n_gpus": world, "steps": nsteps, "warmup": min(args.warmup, 2), "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "PM3 SCF+forces, one (H2O)_%d cluster (%d atoms, %d packed orbitals), SP2 eps 1e-5, converger "
                               "[2], scf_eps 1e-6 (BASELINE.json configs[4])" % (nwater, 3 * nwater, m),
                   "sp2": ("tile-sharded over %d GPUs, %s" % (world, "ncclAllGather of the tile records" if os.environ.get('SEQM_SP2_P2P') == '0' else "records stored peer to peer over NVLink by the multiplication kernel")) if world > 1 else "single GPU",
                   "l2": "working set (X 2 x %d MB, integrals %d MB) exceeds L2" % (m * m * 8 // 10 ** 6, 900 * nwater // 500),
                   "scf_iterations": int(st['iterations']) if st.get('iterations') is not None else None,
                   "scf_not_converged": int(out[10].sum())},
        "e2e": {"value": 1.0 / (e2e_ms / nsteps * 1e-3), "unit": "Force calls/s", "h2d_bytes_per_step": 3 * nwater * 24,
                "d2h_bytes_per_step": 3 * nwater * 24 + 8},
        "gpu_launches": launches, "clocks": sampler.summary(), "roofline": roof, "profile_one_step": prof,
        "parity": parity,
        "cpu_baseline": {"value": 1.0 / 592.7, "unit": "Force calls/s", "cores": 8, "kind": "reference",
                         "sample": "the unmodified reference on this very system in the BUILD container (8 cores, 592.7 s, "
                                   "recorded as ref_seconds in tests/golden_big/water500_PM3_sp2_c2.npz); too long to "
                                   "repeat inside a bench run"}})


def run_organics(ctx):
    """Fourth workload (BASELINE.json configs[3]): MNDO SCF + forces over 65,536 mixed organics of 5-30 atoms (library of
    2048 grown molecules x 32 perturbed copies), eigh density.  STRONG scaling: the batch is cut into contiguous molecule
    ranges, one per rank (pyseqm_b200.distributed.shard_batch); every rank runs its shard as a size-bucketed sweep
    (pyseqm_b200.batching.BucketedBatch); one all_gather of the energies ends the step.  One step = the whole sweep."""
    from pyseqm_b200 import _lib, synth
    from pyseqm_b200.basics import Force
    from pyseqm_b200.batching import BucketedBatch
    from pyseqm_b200.distributed import gather_molecules, shard_bounds
    from pyseqm_b200.seqm_functions.constants import Constants
    ntotal = int(os.environ.get('SEQM_BENCH_ORGANICS', '65536'))
    args, rank, world, dev, local_rank = ctx.args, ctx.rank, ctx.world, ctx.dev, ctx.local_rank
    if world > 1:
        import torch.distributed as dist
    L = _lib.lib()
    nsteps = max(1, min(args.steps, int(os.environ.get('SEQM_BENCH_ORGANICS_STEPS', '3'))))
    par = dict(method='MNDO', scf_eps=1.0e-6, scf_converger=[2, 0.0], sp2=[False], elements=[0, 1, 6, 7, 8], learned=[],
               pair_outer_cutoff=1.0e10)
    sp, xyz = synth.mixed_organics_sweep(ntotal, seed=0)
    log("organics: %d molecules generated" % ntotal)
    lo, hi = shard_bounds(ntotal, world)[rank]
    species = torch.as_tensor(sp[lo:hi]).to(dev)
    coords_host = torch.as_tensor(xyz[lo:hi]).pin_memory()
    coords = coords_host.to(dev)
    nloc = hi - lo
    const = Constants().to(dev)
    fm = Force(par).to(dev)
    bb = BucketedBatch(species)
    log("organics: buckets built")
    workers = int(os.environ.get('SEQM_SWEEP_WORKERS', '6'))      # buckets run concurrently (threads + streams); 0: one after the other

    def step(x):
        f, Etot, Hf, nc = bb.run(fm, const, x, workers=workers)
        if world > 1:
            gather_molecules(Etot, ntotal, world)
        return f, Etot, nc

    barrier = ctx.barrier
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    for _ in range(min(args.warmup, 2)):
        f, Etot, nc = step(coords)
    nfail = int(nc.sum())
    barrier()
    sampler.mark_begin()
    l0 = L.seqm_launch_count()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(nsteps):
        step(coords)
    b.record()
    barrier()
    launches = int(L.seqm_launch_count() - l0)
    total_ms = a.elapsed_time(b)
    f_host = torch.empty(nloc, 30, 3).pin_memory()
    e_host = torch.empty(nloc).pin_memory()
    barrier()
    t0 = time.perf_counter()
    for _ in range(nsteps):
        f, Etot, nc = step(coords_host.to(dev, non_blocking=True))
        f_host.copy_(f, non_blocking=True)
        e_host.copy_(Etot, non_blocking=True)
        torch.cuda.synchronize()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    sampler.mark_end()
    if rank == 0:
        sampler.stop()
    t = torch.tensor([total_ms, e2e_ms, float(nfail)], device=dev)
    if world > 1:
        dist.all_reduce(t[:2], op=dist.ReduceOp.MAX)
        nf = t[2:].clone()
        dist.all_reduce(nf)
        t[2] = nf[0]
    total_ms, e2e_ms, nfail = float(t[0]), float(t[1]), int(t[2])
    if rank != 0:
        return None
    L.seqm_profile_enable(1)
    bb.run(fm, const, coords)
    torch.cuda.synchronize()
    prof, tot_prof = profile_report(L)
    top = max(prof, key=lambda k: prof[k]["ms"])
    hbm, peak_src = peaks()
    # integrals + matrices each SCF kernel must move at least once, summed over the buckets
    alg = 0
    for idx, bsp in bb.buckets:
        nx, nh = (bsp > 1).sum(1), (bsp == 1).sum(1)
        m = 4 * nx + nh
        alg += int((nx * (nx - 1) // 2 * 800 + nx * nh * 80 + nh * (nh - 1) // 2 * 8 + 4 * m * m * 8 + (nx + nh) * 40).sum())
    small_ms = prof.get('scf_small', {"ms": 0.0})["ms"]
    roof = {"bound": "hbm", "kernel": top, "share_of_step": prof[top]["share"],
            "achieved": alg / (tot_prof * 1e-3) / 1e9, "peak": hbm, "peak_source": peak_src, "unit": "GB/s",
            "frac": alg / (tot_prof * 1e-3) / 1e9 / hbm, "traffic": None,
            "note": "algorithmic bytes of the whole sweep (integrals once + 4 packed matrices per molecule) over the sum of "
                    "the profiled kernel times of one sweep; the SCF kernels are latency / FP64-issue bound (DESIGN.md), "
                    "one-launch warp SCF %.1f ms of %.1f ms" % (small_ms, tot_prof)}
    cpu_base = None
    if world == 1 and CPU_LEG:      # CPU baselines are an N = 1 item
        log("organics: cpu baseline")
        cpu = CpuPath()
        ns = 512
        cpu_s = cpu.force_seconds(sp[:ns], xyz[:ns], 'MNDO')
        cpu_base = {"value": ns / cpu_s, "unit": "molecules/s", "cores": cpu.cores, "kind": cpu.kind,
                    "sample": "the first %d molecules of the sweep, one Force call" % ns}
    ms_step = total_ms / nsteps
    return ({
        "metric": "SCF molecules/sec", "value": ntotal / (ms_step * 1e-3), "unit": "molecules/s", "n_gpus": world,
        "steps": nsteps, "warmup": min(args.warmup, 2), "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "MNDO SCF+forces sweep, %d mixed organics of 5-30 atoms (2048 grown molecules x perturbed "
                               "copies), eigh, converger [2], scf_eps 1e-6, sharded by molecule over %d GPU(s) "
                               "(BASELINE.json configs[3])" % (ntotal, world),
                   "molecules_per_gpu": nloc, "buckets_rank0": bb.summary(), "concurrent_bucket_workers": workers,
                   "l2": "working set of a sweep (integrals + matrices of %d molecules) exceeds L2" % nloc,
                   "scf_not_converged": nfail},
        "e2e": {"value": ntotal / (e2e_ms / nsteps * 1e-3), "unit": "molecules/s",
                "h2d_bytes_per_step": nloc * 30 * 3 * 8, "d2h_bytes_per_step": nloc * 30 * 3 * 8 + nloc * 8},
        "gpu_launches": launches, "clocks": sampler.summary(), "roofline": roof, "profile_one_step": prof,
        "cpu_baseline": cpu_base})


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=100)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='b200')
    ap.add_argument('--workload', default='all', choices=['all', 'benzene', 'c60xl', 'water', 'organics'])
    args = ap.parse_args()
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    if args.impl == 'reference':
        run_reference(args, rank, world)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (pyseqm_b200 has no CPU path; --impl reference times the CPU reference)")
    ctx = Ctx(args, rank, world, local_rank)
    torch.cuda.set_device(ctx.dev)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group('nccl', device_id=ctx.dev)
    runners = {'benzene': run_gpu, 'c60xl': run_c60xl, 'water': run_water, 'organics': run_organics}
    if args.workload != 'all':
        line = runners[args.workload](ctx)
    else:
        # ONE line: the headline metric (SCF molecules/s, configs[1]) at top level, the XL-BOMD half of the metric
        # (configs[2]) under "xlbomd"; with more than one GPU also the strong-scaling sweep (configs[3]) and the
        # single large system with the SP2 tiles sharded over the ranks (configs[4]), so that a scaling run exercises
        # every multi-GPU form.  SEQM_BENCH_SUB=0 keeps only the headline; =all adds organics and water at N = 1 too.
        sub = os.environ.get('SEQM_BENCH_SUB', 'auto')
        log("benzene")
        line = run_gpu(ctx)
        extra = {}
        if sub != '0':
            log("xlbomd")
            extra['xlbomd'] = run_c60xl(ctx)
            if world > 1 or sub == 'all':
                log("organics")
                extra['organics'] = run_organics(ctx)
                log("water")
                extra['water'] = run_water(ctx)
        log("done")
        if line is not None:
            line.update({k: v for k, v in extra.items() if v is not None})
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0 and line is not None:
        print(json.dumps(line))


if __name__ == '__main__':
    main()
