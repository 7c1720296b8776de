"""Per-kernel timing of the differentiable trace on the bench lens (C2) or the Cooke triplet.
usage: python profiles/recipes/prof_fwd_bwd.py [c2|cooke] [nrays] [once]   ("once": one fwd+bwd per mode, for ncu)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import numpy as np, torch
from optiland_pr_b200 import samples
from optiland_pr_b200.autograd import trace_differentiable
from optiland_pr_b200.prescription import pack_surface_group
from optiland_pr_b200.trace import fused_trace

which = sys.argv[1] if len(sys.argv) > 1 else "c2"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 24
once = len(sys.argv) > 3
dev = torch.device("cuda")
rng = np.random.default_rng(0)
r = torch.from_numpy(rng.uniform(size=n)).to(dev); th = torch.from_numpy(rng.uniform(0, 2 * np.pi, n)).to(dev)
Px, Py = torch.sqrt(r) * torch.cos(th), torch.sqrt(r) * torch.sin(th)
if which == "c2":
    lens = samples.EvenAsphereDoubleGauss(); wls = [0.5876]
    rays = lens.generate_rays(0.0, 0.7, Px, Py, 0.5876)
    ins = rays.components(); ins[7] = None
else:
    lens = samples.CookeTriplet(); wls = [float(w) for w in lens.wavelengths]
    wl = torch.tensor(wls, dtype=torch.float64, device=dev)[torch.arange(n, device=dev) % 3]
    rays = lens.generate_rays(torch.zeros_like(Px), torch.full_like(Px, 0.7), Px, Py, wl)
    ins = rays.components()
ins = [t.detach() if t is not None else None for t in ins]
S = len(lens.surface_group.surfaces) - 1
g = torch.zeros((8, n), dtype=torch.float64, device=dev); g[0] = 1.0; g[1] = 0.5

def tm(fn, k=5):
    if once:
        fn(); torch.cuda.synchronize(); return float("nan")
    fn(); fn(); torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(k): fn()
    e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1) / k

pk_plain = pack_surface_group(lens.surface_group.surfaces, wls, record="none")
fused_trace(pk_plain, ins, fast=True)     # settles the Newton counts
iters = pk_plain.nr_iters()
out_buf = torch.empty((8, n), dtype=torch.float64, device=dev)
t_plain = tm(lambda: fused_trace(pk_plain, ins, fast=True, verify_nr=False, check_status=False, out=out_buf))
print(f"{which} n={n} S={S} K={iters}: fwd (no checkpoints) {t_plain:.3f} ms")
from optiland_pr_b200 import _lib as L
resid = torch.zeros((pk_plain.n_surfaces, L.RESID_STRIDE), dtype=torch.int64, device=dev)
t_ver = tm(lambda: (resid.zero_(), fused_trace(pk_plain, ins, fast=True, check_status=False, out=out_buf, resid_out=resid)))
print(f"  fwd with Newton verification (residual bits): {t_ver:.3f} ms")
for label, record, blocked, ray_grads, noxy in (("records, all grads", "all", False, True, False),
                                                ("blocked, all grads", "none", True, True, False),
                                                ("blocked, presc only", "none", True, False, False),
                                                ("blocked, presc only, no frame xy", "none", True, False, True)):
    pk = pack_surface_group(lens.surface_group.surfaces, wls, record=record)
    pk.set_nr_iters(iters)
    if noxy:
        pk.live_offsets = frozenset(o for lay in pk.layouts for o in range(lay.frames + 2, lay.aper))
    presc = torch.tensor(pk.presc[0], dtype=torch.float64, device=dev, requires_grad=True)
    leaf = [t.detach().requires_grad_(ray_grads and q < 6) if t is not None else None for q, t in enumerate(ins)]

    def fwd_only():
        with torch.no_grad():
            trace_differentiable(pk, presc, leaf, fast=True, blocked=blocked)

    def fwd_bwd():
        presc.grad = None
        out, rec, _ = trace_differentiable(pk, presc, leaf, fast=True, blocked=blocked)
        out.backward(g)

    a, b = tm(fwd_only), tm(fwd_bwd)
    print(f"  {label:34s}: fwd {a:.3f} ms, fwd+bwd {b:.3f} ms, bwd {b - a:.3f} ms  ({n * S / b / 1e6:.1f} G rs/s)")
    del presc, leaf
    torch.cuda.empty_cache()
