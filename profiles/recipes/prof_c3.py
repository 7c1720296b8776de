"""BASELINE C3 (three-mirror polynomial freeform, interpreter kernel) for timing / ncu. usage: prof_c3.py [n] [strict]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import optiland_pr_b200 as ol
from optiland_pr_b200 import samples
from optiland_pr_b200.trace import fused_trace
ol.backend.set_device("cuda:0"); ol.backend.set_precision("float64")
n3 = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 24
fast = not (len(sys.argv) > 2 and sys.argv[2] == "strict")
lens = samples.ThreeMirrorAnastigmat(); wl = lens.primary_wavelength
rng = np.random.default_rng(0)
r_, th = rng.uniform(size=n3), rng.uniform(0, 2 * np.pi, size=n3)
Px = torch.from_numpy(np.sqrt(r_) * np.cos(th)).cuda(); Py = torch.from_numpy(np.sqrt(r_) * np.sin(th)).cuda()
f0 = lens.get_field_coords()[0]
rays = lens.launch_rays(np.array([f0[0]]), np.array([f0[1]]), Px, Py, wl)
ins = rays.components(); ins[7] = None
lens.surface_group.record = "none"
packed = lens.surface_group.pack([float(wl)])
o = torch.empty((8, n3), dtype=torch.float64, device="cuda")
res = fused_trace(packed, ins, out=o, fast=fast)
print("nr_iters", res.nr_iters, "nan frac", float(o[0].isnan().float().mean()))
for _ in range(3): fused_trace(packed, ins, out=o, fast=fast, verify_nr=False, check_status=False)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5): fused_trace(packed, ins, out=o, fast=fast, verify_nr=False, check_status=False)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print(f"n={n3} fast={fast}: {ms:.3f} ms, {n3*4/ms/1e6:.1f} G ray-surfaces/s")
from optiland_pr_b200 import _lib as L
resid = torch.zeros((packed.n_surfaces, L.RESID_STRIDE), dtype=torch.int64, device="cuda")
st = torch.zeros(1, dtype=torch.int32, device="cuda")
for _ in range(3): fused_trace(packed, ins, out=o, fast=fast, verify_nr=False, check_status=False, resid_out=resid, status_t=st)
torch.cuda.synchronize()
e0.record()
for _ in range(5): fused_trace(packed, ins, out=o, fast=fast, verify_nr=False, check_status=False, resid_out=resid, status_t=st)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print(f"n={n3} fast={fast} WITH residual report: {ms:.3f} ms, {n3*4/ms/1e6:.1f} G ray-surfaces/s")
