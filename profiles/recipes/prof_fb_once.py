"""One forward (blocked checkpoints) + one backward (reduced-algebra adjoint, prescription gradients only, no frame
x / y) of the C2 lens between cudaProfilerStart / Stop, for `ncu --profile-from-start off`.
usage: python profiles/recipes/prof_fb_once.py [c2|cooke] [nrays]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import numpy as np, torch
from optiland_pr_b200 import samples
from optiland_pr_b200.autograd import trace_differentiable
from optiland_pr_b200.optic import random_pupil_device
from optiland_pr_b200.prescription import pack_surface_group
from optiland_pr_b200.trace import fused_trace
import optiland_pr_b200 as ol
which = sys.argv[1] if len(sys.argv) > 1 else "c2"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 23
ol.backend.set_device("cuda:0"); ol.backend.set_precision("float64")
dev = torch.device("cuda")
Px, Py = random_pupil_device(n, 0, dev)
if which == "c2":
    lens = samples.EvenAsphereDoubleGauss(); wls = [0.5876]
    rays = lens.generate_rays(torch.zeros_like(Px), torch.full_like(Px, 0.7), Px, Py, 0.5876)
    ins = rays.components(); ins[7] = None
else:
    lens = samples.CookeTriplet(); wls = [float(w) for w in lens.wavelengths]
    wl = torch.tensor(wls, dtype=torch.float64, device=dev)[torch.arange(n, device=dev) % 3]
    rays = lens.generate_rays(torch.zeros_like(Px), torch.full_like(Px, 0.7), Px, Py, wl)
    ins = rays.components()
ins = [t.detach() if t is not None else None for t in ins]
pk = pack_surface_group(lens.surface_group.surfaces, wls, record="none")
fused_trace(pk, ins, fast=True)
pk.live_offsets = frozenset(o for lay in pk.layouts for o in range(lay.frames + 2, lay.aper))
presc = torch.tensor(pk.presc[0], dtype=torch.float64, device=dev, requires_grad=True)
g = torch.zeros((8, n), dtype=torch.float64, device=dev); g[0] = 1.0; g[1] = 0.5

def fb():
    presc.grad = None
    out, _, _ = trace_differentiable(pk, presc, ins, fast=True, blocked=True)
    out.backward(g)

fb(); torch.cuda.synchronize()
torch.cuda.profiler.start()
fb(); torch.cuda.synchronize()
torch.cuda.profiler.stop()
print(f"prof_fb_once ok: {which}, {n} rays, K={pk.nr_iters()}")
