"""fwd (all surfaces recorded) + bwd on the bench lens, for ncu. usage: python scratch/prof_bwd.py [nrays]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from optiland_pr_b200 import samples
from optiland_pr_b200.autograd import trace_differentiable
from optiland_pr_b200.prescription import pack_surface_group
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 22
dev = torch.device("cuda")
lens = samples.EvenAsphereDoubleGauss()
rng = np.random.default_rng(0)
r = torch.from_numpy(rng.uniform(size=n)).to(dev); th = torch.from_numpy(rng.uniform(0, 2*np.pi, n)).to(dev)
rays = lens.generate_rays(0.0, 0.7, torch.sqrt(r)*torch.cos(th), torch.sqrt(r)*torch.sin(th), 0.5876)
ins = rays.components(); ins[7] = None
pk = pack_surface_group(lens.surface_group.surfaces, [0.5876], record="all")
presc = torch.tensor(pk.presc[0], dtype=torch.float64, device=dev, requires_grad=True)
leaf = [t.detach().requires_grad_(q < 6) if t is not None else None for q, t in enumerate(ins)]
g = torch.zeros((8, n), dtype=torch.float64, device=dev); g[0] = 1.0; g[1] = 0.5
for it in range(4):
    presc.grad = None
    out, rec, _ = trace_differentiable(pk, presc, leaf)
    out.backward(g)
torch.cuda.synchronize()
print("ok", float(presc.grad.abs().sum()))
def tm(fn, k=5):
    fn(); torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(k): fn()
    e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1) / k
def fwd_only():
    with torch.no_grad():
        trace_differentiable(pk, presc, leaf, fast=True)
def fwd_bwd():
    presc.grad = None
    out, rec, _ = trace_differentiable(pk, presc, leaf, fast=True)
    out.backward(g)
a, b = tm(fwd_only), tm(fwd_bwd)
print(f"n={n}: fwd(record all) {a:.3f} ms, fwd+bwd {b:.3f} ms, bwd {b-a:.3f} ms  ({n*12/b/1e6:.1f} G rs/s fwd+bwd)")
pkb = pack_surface_group(lens.surface_group.surfaces, [0.5876], record="none"); pkb.set_nr_iters(pk.nr_iters())
def fwd_only_b():
    with torch.no_grad():
        trace_differentiable(pkb, presc, leaf, fast=True, blocked=True)
def fwd_bwd_b():
    presc.grad = None
    out, rec, _ = trace_differentiable(pkb, presc, leaf, fast=True, blocked=True)
    out.backward(g)
if len(sys.argv) > 2:
    a, b = tm(fwd_only_b), tm(fwd_bwd_b)
    print(f"n={n}: BLOCKED fwd {a:.3f} ms, fwd+bwd {b:.3f} ms, bwd {b-a:.3f} ms  ({n*12/b/1e6:.1f} G rs/s fwd+bwd)")
