"""The bench's dominant launch (olrt_trace_pupil_f64: ray launch + trace + final state + spot sums) on the bench lens,
for ncu.  usage: python profiles/recipes/prof_fx.py [rays_per_field]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import optiland_pr_b200 as ol
from optiland_pr_b200 import _lib, samples
from optiland_pr_b200.trace import fused_trace_pupil
npf = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 24
dev = torch.device("cuda")
ol.backend.set_device("cuda:0"); ol.backend.set_precision("float64")
lens = samples.EvenAsphereDoubleGauss(); lens.surface_group.record = "none"
rng = np.random.default_rng(1)
r = torch.from_numpy(rng.uniform(size=npf)).to(dev); th = torch.from_numpy(rng.uniform(0, 2*np.pi, npf)).to(dev)
Px, Py = torch.sqrt(r)*torch.cos(th), torch.sqrt(r)*torch.sin(th)
Hx, Hy = np.array([0.0, 0.0, 0.0]), np.array([0.0, 10.0 / 14.0, 1.0])
packed = lens.surface_group.pack([0.5876])
packed.set_nr_iters([1, 0, 0, 0, 2, 0, 1, 0, 0, 0, 1, 0])   # the counts bench.py settles on
fxy_np, z_start = lens._field_offsets(Hx, Hy)
fxy = torch.from_numpy(fxy_np).to(dev)
kw = dict(EPD=lens._EPD, EPL=lens._EPL, z_start=z_start, wavelength=0.5876, dtype=torch.float64)
n = 3 * npf
out = torch.empty((8, n), dtype=torch.float64, device=dev)
sums = torch.zeros((3, _lib.SPOT_SUMS), dtype=torch.float64, device=dev)
shift = torch.zeros((3, 2), dtype=torch.float64, device=dev)
# per-field Newton counts and residual reports, as in the timed step of bench.py (trace_fx_kernel<double, 25>)
kf = torch.tensor([[1, 0, 0, 0, 1, 0, 1, 0, 0, 0, 1, 0]] * 2 + [[1, 0, 0, 0, 2, 0, 1, 0, 0, 0, 1, 0]], dtype=torch.int32, device=dev)
gres = torch.zeros((3, packed.n_surfaces, _lib.GROUP_RESID), dtype=torch.int64, device=dev)
kw.update(group_k=kf, group_resid=gres)
for _ in range(3):
    fused_trace_pupil(packed, Px, Py, fxy, out=out, spot=sums, spot_shift=shift, **kw)
torch.cuda.synchronize()
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    fused_trace_pupil(packed, Px, Py, fxy, out=out, spot=sums, spot_shift=shift, **kw)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print(f"n={n}: {ms:.3f} ms per launch, {n*12/ms/1e6:.1f} G ray-surfaces/s")
