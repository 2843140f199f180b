#!/usr/bin/env python
"""Benchmark of the batched SCF hot path (BASELINE.json config 2): AM1 SCF + forces for a batch of 1024
perturbed benzene geometries (12 atoms, 30 orbitals), eigh density path; metric = SCF molecules / second.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload benzene|c60xl|water|organics]

One step = one `Force` call (Parser + pair pass + SCF + force adjoint) over the whole batch.  `value` is
timed with the coordinates resident in HBM; `e2e` includes, every step, the host->device copy of the
coordinates from pinned memory and the device->host copy of forces and energies.  N > 1: one process per
GPU (torchrun), every rank runs its own 1024-molecule shard (weak scaling), no collective in the loop,
one all_gather of the energies at the end of every step.
"""
import argparse
import datetime
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
torch.set_default_dtype(torch.float64)

NMOL = 1024
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


def run_reference(args, rank, world):
    """The reference's algorithm on the host cores (oracle port: the reference itself is a Python package that
    does not exist on the GPU box).  Each step = a bounded sample of the same workload."""
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, 'oracle'))
    import seqm_oracle as orc
    from pyseqm_b200 import synth
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    nsample = 512
    sp, xyz = synth.benzene(nsample, seed=0)
    species, coords = torch.as_tensor(sp), torch.as_tensor(xyz)
    times = []
    for it in range(args.warmup + args.steps):
        t0 = time.time()
        orc.run(species, coords, method='AM1', scf_eps=1e-6, converger=(2, 0.0))
        if it >= args.warmup:
            times.append(time.time() - t0)
    ms = 1e3 * sum(times) / len(times)
    val = nsample / (ms * 1e-3)
    sample = "%d of the 1024 perturbed benzenes per step (same generator, seed 0)" % nsample
    print(json.dumps({
        "impl": "reference", "metric": "SCF molecules/sec", "value": val, "unit": "molecules/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "AM1 SCF+forces, perturbed benzene (12 atoms, 30 orbitals), eigh, converger [2], "
                               "scf_eps 1e-6", "sample": sample},
        "cpu_baseline": {"value": val, "unit": "molecules/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "molecules/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


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


def run_gpu(args, rank, world, local_rank):
    import ctypes
    from pyseqm_b200 import _lib, _ops, synth
    from pyseqm_b200.basics import Force
    from pyseqm_b200.seqm_functions.constants import Constants
    dev = torch.device('cuda', local_rank)
    torch.cuda.set_device(dev)
    if world > 1:
        import torch.distributed as dist
        from pyseqm_b200.distributed import gather_molecules
        dist.init_process_group('nccl', device_id=dev)
    L = _lib.lib()
    sp, xyz = synth.benzene(NMOL, seed=rank)
    species = torch.as_tensor(sp).to(dev)
    coords_host = torch.as_tensor(xyz).pin_memory()
    coords = coords_host.to(dev)
    const = Constants().to(dev)
    force_mod = Force(dict(SEQM_PAR)).to(dev)
    flush_buf = torch.empty(192 * 1024 * 1024, dtype=torch.uint8, device=dev)     # > 126 MB of L2

    def flush():
        flush_buf.fill_(1)

    def step(x):
        out = force_mod(const, x, species)
        if world > 1:
            gather_molecules(out[2], NMOL * world, world)     # energies of every shard, global molecule order
        return out

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    for _ in range(args.warmup):
        out = step(coords.clone())
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
    barrier()
    launches = int(L.seqm_launch_count() - l0)
    total_ms = sum(a.elapsed_time(b) for a, b in evs)
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
    e2e_ms = (time.perf_counter() - t0) * 1e3
    sampler.mark_end()
    if rank == 0:
        sampler.stop()
    t = torch.tensor([total_ms, e2e_ms], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, e2e_ms = float(t[0]), float(t[1])
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    # ---- per-kernel profile of one step, and the kernels timed alone on the converged state ----
    L.seqm_profile_enable(1)
    force_mod(const, coords.clone(), species)        # no collective here: the other ranks have finished
    buf = ctypes.create_string_buffer(1 << 16)
    L.seqm_profile_report(buf, len(buf))
    L.seqm_profile_enable(0)
    prof = {}
    for line in buf.value.decode().splitlines():
        nm, ms, cnt = line.rsplit(' ', 2)
        prof[nm] = {"ms": float(ms), "launch_groups": int(cnt)}
    tot_prof = sum(v["ms"] for v in prof.values()) or 1.0
    for v in prof.values():
        v["share"] = v["ms"] / tot_prof
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
        roof.update({"traffic": 198338048 + 125199872,
                     "traffic_source": "ncu dram__bytes_read.sum + dram__bytes_write.sum, one launch (cold-cache replay), "
                                       "profiles/r01p_scf_small_ncu.txt",
                     "limiter": "latency of one warp per molecule (ncu: sm__pipe_fp64_cycles_active 19.5 %, "
                                "smsp__issue_active 26.0 %, stalls short_scoreboard / wait / long_scoreboard)",
                     "algorithmic_tflops": flops / (kernels[top]["ms_per_launch"] * 1e-3) / 1e12,
                     "algorithmic_flops_note": "(9 m^3 + 2 m^2 nocc) per eigensolve x (SCF iterations + 1) x molecules"})
    # ---- CPU baseline: the oracle port on a bounded sample, all host cores ----
    sys.path.insert(0, os.path.join(ROOT, 'oracle'))
    import seqm_oracle as orc
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    ns, reps = NMOL, 4
    t0 = time.time()
    for _ in range(reps):
        orc.run(torch.as_tensor(sp), torch.as_tensor(xyz), method='AM1', scf_eps=1e-6, converger=(2, 0.0))
    cpu_s = (time.time() - t0) / reps
    ms_step = total_ms / args.steps
    nm_total = NMOL * world
    print(json.dumps({
        "metric": "SCF molecules/sec", "value": nm_total / (ms_step * 1e-3), "unit": "molecules/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "AM1 SCF+forces, 1024 perturbed benzenes per GPU (12 atoms, 30 orbitals), eigh, "
                               "converger [2], scf_eps 1e-6 (BASELINE.json configs[1])",
                   "molecules_per_gpu": NMOL, "l2": "192 MiB buffer written between timed steps",
                   "scf_not_converged": nconv_fail, "scf_iterations": int(st['iterations'])},
        "e2e": {"value": nm_total / (e2e_ms / args.steps * 1e-3), "unit": "molecules/s",
                "h2d_bytes_per_step": NMOL * 12 * 3 * 8, "d2h_bytes_per_step": NMOL * 12 * 3 * 8 + NMOL * 8},
        "gpu_launches": launches, "clocks": sampler.summary(), "roofline": roof, "kernels": kernels,
        "profile_one_step": prof,
        "cpu_baseline": {"value": ns / cpu_s, "unit": "molecules/s", "cores": cores, "kind": "port",
                         "sample": "the %d benzenes of rank 0, %d Force-equivalent calls of the oracle" % (ns, reps)}}))
    if world > 1:
        dist.destroy_process_group()


def run_c60xl(args, rank, world, local_rank):
    """Secondary workload (BASELINE.json configs[2]): XL-BOMD (k=5) of 256 C60 per GPU, PM3, SP2 density
    (eps 1e-5); one step = one `XL_BOMD.one_step` of the whole batch (pair pass, Fock build with the dynamical
    density, one SP2 purification, shadow energy, force adjoint, Verlet).  metric: batch MD steps per second."""
    import ctypes
    from pyseqm_b200 import _lib, synth
    from pyseqm_b200.XLBOMD import XL_BOMD
    from pyseqm_b200.seqm_functions.constants import Constants
    nmol = int(os.environ.get('SEQM_BENCH_C60', '256'))
    dev = torch.device('cuda', local_rank)
    torch.cuda.set_device(dev)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group('nccl', device_id=dev)
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

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    for _ in range(args.warmup):
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
    launches = int(L.seqm_launch_count() - l0)
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
    t = torch.tensor([total_ms, e2e_ms], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, e2e_ms = float(t[0]), float(t[1])
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    L.seqm_profile_enable(1)
    step()
    buf = ctypes.create_string_buffer(1 << 16)
    L.seqm_profile_report(buf, len(buf))
    L.seqm_profile_enable(0)
    prof = {}
    for line in buf.value.decode().splitlines():
        nm, ms, cnt = line.rsplit(' ', 2)
        prof[nm] = {"ms": float(ms), "launch_groups": int(cnt)}
    tot_prof = sum(v["ms"] for v in prof.values()) or 1.0
    for v in prof.values():
        v["share"] = v["ms"] / tot_prof
    # FP64 GEMM peak of this box, measured here (cuBLAS DGEMM through torch.matmul, 4096^3, best of 5): denominator
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
    fp64_peak = 2.0 * 4096 ** 3 / (best * 1e-3) / 1e12
    del A_, B_
    m = 240
    sp = prof.get('sp2_step', {"ms": 0.0, "launch_groups": 1, "share": 0.0})
    nstep_launch = 2 * max(sp["launch_groups"], 1)          # two purification steps per profiled group
    flops = 2.0 * m ** 3 * nmol                              # 2 m^3 per multiplication and molecule
    per_launch = sp["ms"] / nstep_launch
    T_ = (m + 63) // 64
    roof = {"bound": "tensor", "kernel": "sp2_step_kernel (X^2 on the upper block triangle, DMMA m8n8k4 f64, fused trace "
                                         "rule / update / transpose epilogue)", "share_of_step": sp["share"],
            "achieved": flops / (per_launch * 1e-3) / 1e12 if per_launch else None, "peak": fp64_peak,
            "peak_source": "cuBLAS DGEMM 4096^3 measured in this run (MEASURED_PEAKS.json holds no FP64 number)",
            "unit": "TFLOP/s", "frac": flops / (per_launch * 1e-3) / 1e12 / fp64_peak if per_launch else None,
            "traffic": (118043904 + 79299328) if nmol == 256 else None,
            "traffic_source": "ncu dram bytes of one sp2_step_kernel launch, profiles/r01p_c60_ncu.txt (top stall: "
                              "math_pipe_throttle, tensor pipe = DMMA)",
            "note": "algorithmic flops 2 m^3 per product; the kernel executes %d of %d tiles (symmetry), i.e. %.3f of "
                    "them; the last launch of the loop may be a no-op" % (T_ * (T_ + 1) // 2, T_ * T_,
                                                                        (T_ + 1) / (2.0 * T_))}
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
    print(json.dumps({
        "metric": "XL-BOMD steps/sec", "value": world / (ms_step * 1e-3), "unit": "batch steps/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "XL-BOMD k=5, %d C60 per GPU (60 atoms, 240 orbitals), PM3, SP2 eps 1e-5 "
                               "(BASELINE.json configs[2])" % nmol, "molecules_per_gpu": nmol,
                   "l2": "working set per step (P, F, D, w, ring: > 1 GB) exceeds L2"},
        "e2e": {"value": world / (e2e_ms / args.steps * 1e-3), "unit": "batch steps/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": nmol * 60 * 3 * 8 + nmol * 8,
                "note": "MD state lives on the device (as in the reference); coordinates and Hf go to pinned host "
                        "memory every step"},
        "gpu_launches": launches, "clocks": sampler.summary(), "roofline": roof, "kernels": kernels,
        "profile_one_step": prof, "cpu_baseline": None}))
    if world > 1:
        dist.destroy_process_group()


def run_water(args, rank, world, local_rank):
    """Third workload (BASELINE.json configs[4]): PM3 SCF + forces of ONE (H2O)_500 cluster (1500 atoms, 3000 packed
    orbitals), SP2 density (eps 1e-5).  One step = one `Force` call.  N > 1: STRONG scaling - every rank holds the
    whole system (pair pass, Fock build and gradient replicated, < 10 % of the step), the SP2 multiplications are
    tile-sharded over the ranks with ncclAllGather (pyseqm_b200.distributed.enable_sp2_sharding)."""
    import ctypes
    from pyseqm_b200 import _lib, synth
    from pyseqm_b200.basics import Force
    from pyseqm_b200.seqm_functions.constants import Constants
    nwater = int(os.environ.get('SEQM_BENCH_WATER', '500'))
    dev = torch.device('cuda', local_rank)
    torch.cuda.set_device(dev)
    if world > 1:
        import torch.distributed as dist
        from pyseqm_b200.distributed import disable_sp2_sharding, enable_sp2_sharding
        dist.init_process_group('nccl', device_id=dev)
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

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    for _ in range(args.warmup):
        out = fm(const, coords.clone(), species)
    barrier()
    sampler.mark_begin()
    l0 = L.seqm_launch_count()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(args.steps):
        out = fm(const, coords.clone(), species)
    b.record()
    barrier()
    launches = int(L.seqm_launch_count() - l0)
    total_ms = a.elapsed_time(b)
    f_host = torch.empty(1, 3 * nwater, 3).pin_memory()
    e_host = torch.empty(1).pin_memory()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        o = fm(const, coords_host.to(dev, non_blocking=True), species)
        f_host.copy_(o[0], non_blocking=True)
        e_host.copy_(o[2].detach(), non_blocking=True)
        torch.cuda.synchronize()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    sampler.mark_end()
    if rank == 0:
        sampler.stop()
    t = torch.tensor([total_ms, e2e_ms], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, e2e_ms = float(t[0]), float(t[1])
    # per-kernel profile of one more call: collective inside, so every rank runs it; rank 0 reports
    L.seqm_profile_enable(1 if rank == 0 else 0)
    fm(const, coords.clone(), species)
    torch.cuda.synchronize()
    st = fm.energy.last_state
    if world > 1:
        dist.barrier()
        disable_sp2_sharding()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    buf = ctypes.create_string_buffer(1 << 16)
    L.seqm_profile_report(buf, len(buf))
    L.seqm_profile_enable(0)
    prof = {}
    for line in buf.value.decode().splitlines():
        nm, ms, cnt = line.rsplit(' ', 2)
        prof[nm] = {"ms": float(ms), "launch_groups": int(cnt)}
    tot_prof = sum(v["ms"] for v in prof.values()) or 1.0
    for v in prof.values():
        v["share"] = v["ms"] / tot_prof
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
    fp64_peak = 2.0 * 4096 ** 3 / (best * 1e-3) / 1e12
    del A_, B_
    m = int(st['topo'].mmax)
    spp = prof.get('sp2_step', {"ms": 0.0, "launch_groups": 1, "share": 0.0})
    nstep = 2 * max(spp["launch_groups"], 1)                  # two purification steps per profiled group (upper bound)
    flops = 2.0 * m ** 3                                     # 2 m^3 per multiplication
    per_step = spp["ms"] / nstep
    T_ = (m + 63) // 64
    ach = flops / (per_step * 1e-3) / 1e12 if per_step else None
    roof = {"bound": "tensor", "kernel": "sp2_step_kernel%s" % ("<sharded> + ncclAllGather + sp2_unpack_kernel" if world > 1 else ""),
            "share_of_step": spp["share"], "achieved": ach, "peak": fp64_peak * world,
            "peak_source": "cuBLAS DGEMM 4096^3 measured in this run on rank 0, times the number of GPUs",
            "unit": "TFLOP/s", "frac": ach / (fp64_peak * world) if ach else None,
            "traffic": (181162752 + 51730432) if (world == 1 and nwater == 500) else None,
            "traffic_source": "ncu dram bytes of one single-GPU sp2_step_kernel launch, profiles/r01p_sp2_step_water_ncu.txt "
                              "(sm__pipe_tensor_cycles_active 61 %, top stall math_pipe_throttle)",
            "note": "2 m^3 flops per purification step over the time of one step (multiplication, all-gather and "
                    "unpack) on all GPUs; %d of %d tiles executed (symmetry)" % (T_ * (T_ + 1) // 2, T_ * T_)}
    ms_step = total_ms / args.steps
    print(json.dumps({
        "metric": "SCF systems/sec (one 1500-atom system, SCF + forces)", "value": 1.0 / (ms_step * 1e-3), "unit": "Force calls/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "PM3 SCF+forces, one (H2O)_%d cluster (%d atoms, %d packed orbitals), SP2 eps 1e-5, converger "
                               "[2], scf_eps 1e-6 (BASELINE.json configs[4])" % (nwater, 3 * nwater, m),
                   "sp2": ("tile-sharded over %d GPUs, %s" % (world, "ncclAllGather of the tile records" if os.environ.get('SEQM_SP2_P2P') == '0' else "records stored peer to peer over NVLink by the multiplication kernel")) if world > 1 else "single GPU",
                   "l2": "working set (X 2 x %d MB, integrals %d MB) exceeds L2" % (m * m * 8 // 10 ** 6, 900 * nwater // 500),
                   "scf_iterations": int(st['iterations']) if st.get('iterations') is not None else None,
                   "scf_not_converged": int(out[10].sum())},
        "e2e": {"value": 1.0 / (e2e_ms / args.steps * 1e-3), "unit": "Force calls/s", "h2d_bytes_per_step": 3 * nwater * 24,
                "d2h_bytes_per_step": 3 * nwater * 24 + 8},
        "gpu_launches": launches, "clocks": sampler.summary(), "roofline": roof, "profile_one_step": prof,
        "cpu_baseline": None}))
    if world > 1:
        dist.destroy_process_group()


def run_organics(args, rank, world, local_rank):
    """Fourth workload (BASELINE.json configs[3]): MNDO SCF + forces over 65,536 mixed organics of 5-30 atoms (library of
    2048 grown molecules x 32 perturbed copies), eigh density.  STRONG scaling: the batch is cut into contiguous molecule
    ranges, one per rank (pyseqm_b200.distributed.shard_batch); every rank runs its shard as a size-bucketed sweep
    (pyseqm_b200.batching.BucketedBatch); one all_gather of the energies ends the step.  One step = the whole sweep."""
    import ctypes
    from pyseqm_b200 import _lib, synth
    from pyseqm_b200.basics import Force
    from pyseqm_b200.batching import BucketedBatch
    from pyseqm_b200.distributed import gather_molecules, shard_bounds
    from pyseqm_b200.seqm_functions.constants import Constants
    ntotal = int(os.environ.get('SEQM_BENCH_ORGANICS', '65536'))
    dev = torch.device('cuda', local_rank)
    torch.cuda.set_device(dev)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group('nccl', device_id=dev)
    L = _lib.lib()
    par = dict(method='MNDO', scf_eps=1.0e-6, scf_converger=[2, 0.0], sp2=[False], elements=[0, 1, 6, 7, 8], learned=[],
               pair_outer_cutoff=1.0e10)
    sp, xyz = synth.mixed_organics_sweep(ntotal, seed=0)
    lo, hi = shard_bounds(ntotal, world)[rank]
    species = torch.as_tensor(sp[lo:hi]).to(dev)
    coords_host = torch.as_tensor(xyz[lo:hi]).pin_memory()
    coords = coords_host.to(dev)
    nloc = hi - lo
    const = Constants().to(dev)
    fm = Force(par).to(dev)
    bb = BucketedBatch(species)
    workers = int(os.environ.get('SEQM_SWEEP_WORKERS', '6'))      # buckets run concurrently (threads + streams); 0: one after the other

    def step(x):
        f, Etot, Hf, nc = bb.run(fm, const, x, workers=workers)
        if world > 1:
            gather_molecules(Etot, ntotal, world)
        return f, Etot, nc

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    for _ in range(args.warmup):
        f, Etot, nc = step(coords)
    nfail = int(nc.sum())
    barrier()
    sampler.mark_begin()
    l0 = L.seqm_launch_count()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(args.steps):
        step(coords)
    b.record()
    barrier()
    launches = int(L.seqm_launch_count() - l0)
    total_ms = a.elapsed_time(b)
    f_host = torch.empty(nloc, 30, 3).pin_memory()
    e_host = torch.empty(nloc).pin_memory()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
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
        if world > 1:
            dist.destroy_process_group()
        return
    L.seqm_profile_enable(1)
    bb.run(fm, const, coords)
    torch.cuda.synchronize()
    buf = ctypes.create_string_buffer(1 << 16)
    L.seqm_profile_report(buf, len(buf))
    L.seqm_profile_enable(0)
    prof = {}
    for line in buf.value.decode().splitlines():
        nm, ms, cnt = line.rsplit(' ', 2)
        prof[nm] = {"ms": float(ms), "launch_groups": int(cnt)}
    tot_prof = sum(v["ms"] for v in prof.values()) or 1.0
    for v in prof.values():
        v["share"] = v["ms"] / tot_prof
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
    sys.path.insert(0, os.path.join(ROOT, 'oracle'))
    import seqm_oracle as orc
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    ns = 512
    t0 = time.time()
    orc.run(torch.as_tensor(sp[:ns]), torch.as_tensor(xyz[:ns]), method='MNDO', scf_eps=1e-6, converger=(2, 0.0))
    cpu_s = time.time() - t0
    ms_step = total_ms / args.steps
    print(json.dumps({
        "metric": "SCF molecules/sec", "value": ntotal / (ms_step * 1e-3), "unit": "molecules/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "MNDO SCF+forces sweep, %d mixed organics of 5-30 atoms (2048 grown molecules x perturbed "
                               "copies), eigh, converger [2], scf_eps 1e-6, sharded by molecule over %d GPU(s) "
                               "(BASELINE.json configs[3])" % (ntotal, world),
                   "molecules_per_gpu": nloc, "buckets_rank0": bb.summary(), "concurrent_bucket_workers": workers,
                   "l2": "working set of a sweep (integrals + matrices of %d molecules) exceeds L2" % nloc,
                   "scf_not_converged": nfail},
        "e2e": {"value": ntotal / (e2e_ms / args.steps * 1e-3), "unit": "molecules/s",
                "h2d_bytes_per_step": nloc * 30 * 3 * 8, "d2h_bytes_per_step": nloc * 30 * 3 * 8 + nloc * 8},
        "gpu_launches": launches, "clocks": sampler.summary(), "roofline": roof, "profile_one_step": prof,
        "cpu_baseline": {"value": ns / cpu_s, "unit": "molecules/s", "cores": cores, "kind": "port",
                         "sample": "the first %d molecules of the sweep, one Force-equivalent call of the oracle" % ns}}))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=10)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200')
    ap.add_argument('--workload', default='benzene', choices=['benzene', 'c60xl', 'water', 'organics'])
    args = ap.parse_args()
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    if args.impl == 'reference':
        run_reference(args, rank, world)
    elif args.workload == 'c60xl':
        run_c60xl(args, rank, world, local_rank)
    elif args.workload == 'water':
        run_water(args, rank, world, local_rank)
    elif args.workload == 'organics':
        run_organics(args, rank, world, local_rank)
    else:
        run_gpu(args, rank, world, local_rank)


if __name__ == '__main__':
    main()
