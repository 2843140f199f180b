"""Experiment: one Force call (config 2) captured in a CUDA graph (torch.cuda.graph) and replayed."""
import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
torch.set_default_dtype(torch.float64)
from pyseqm_b200 import synth
from pyseqm_b200.basics import Force
from pyseqm_b200.seqm_functions.constants import Constants
dev = torch.device('cuda:0')
par = dict(method='AM1', scf_eps=1.0e-6, scf_converger=[2, 0.0], sp2=[False], elements=[0, 1, 6], learned=[],
           pair_outer_cutoff=1.0e10, defer_scf_warning=True)
sp, xyz = synth.benzene(1024, seed=0)
species, coords = torch.as_tensor(sp).to(dev), torch.as_tensor(xyz).to(dev)
const = Constants().to(dev)
f = Force(par).to(dev)
ref = f(const, coords.clone(), species)
f.flush_status()
torch.cuda.synchronize()
x_static = coords.clone()
s = torch.cuda.Stream()
s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    for _ in range(3):
        x_static.grad = None
        out = f(const, x_static, species)
    f.flush_status()
torch.cuda.current_stream().wait_stream(s)
torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
x_static.grad = None
with torch.cuda.graph(g):
    out = f(const, x_static, species)
torch.cuda.synchronize()
f._pending.clear()
x_static.data.copy_(coords + 0.01)
g.replay()
torch.cuda.synchronize()
chk = f(const, (coords + 0.01).clone(), species)
torch.cuda.synchronize()
print("graph vs eager: dE %.3e dF %.3e" % (float((out[2] - chk[2]).abs().max()), float((out[0] - chk[0]).abs().max())))
for name, fn in (("eager", lambda: f(const, coords.clone(), species)), ("graph", lambda: g.replay())):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20):
        fn()
    b.record()
    torch.cuda.synchronize()
    print("%s: %.3f ms/step device, %.3f ms/step wall" % (name, a.elapsed_time(b) / 20, (time.perf_counter() - t0) * 50))
