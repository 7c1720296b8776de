"""The fused pupil kernel on the bench batch once without verification (MODE 8) and once with per-field counts and bit
reports (MODE 25), between cudaProfilerStart / Stop, for `ncu --profile-from-start off -k regex:trace_fx_kernel` and a
per-line diff of the two launches (profiles/recipes/ncu_lines.py).  usage: prof_fx_modes.py [rays_per_field]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import numpy as np, torch
import optiland_pr_b200 as ol
from optiland_pr_b200 import _lib, samples, trace as T
from optiland_pr_b200.optic import random_pupil_device
npf = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 22
dev = torch.device("cuda")
ol.backend.set_device("cuda:0"); ol.backend.set_precision("float64")
lens = samples.EvenAsphereDoubleGauss(); lens.surface_group.record = "none"
Hx, Hy = np.zeros(3), np.array([0.0, 10 / 14, 1.0])
Px, Py = random_pupil_device(npf, 1, dev)
packed = lens.surface_group.pack([0.5876])
fxy, z0 = lens._field_offsets(Hx, Hy)
fxy = torch.from_numpy(fxy).to(dev)
kw = dict(EPD=lens._EPD, EPL=lens._EPL, z_start=z0, wavelength=0.5876, dtype=torch.float64)
out = torch.empty((8, 3 * npf), dtype=torch.float64, device=dev)
sums = torch.zeros((3, _lib.SPOT_SUMS), dtype=torch.float64, device=dev)
shift = torch.zeros((3, 2), dtype=torch.float64, device=dev)
S = packed.n_surfaces
kmax = [1, 0, 0, 0, 2, 0, 1, 0, 0, 0, 1, 0]
packed.set_nr_iters(kmax)
gres = torch.zeros((3, S, _lib.GROUP_RESID), dtype=torch.int64, device=dev)
k_own = torch.tensor([[1, 0, 0, 0, 1, 0, 1, 0, 0, 0, 1, 0]] * 2 + [kmax], dtype=torch.int32, device=dev)
plain = lambda: T.fused_trace_pupil(packed, Px, Py, fxy, out=out, spot=sums, spot_shift=shift, **kw)
verified = lambda: T.fused_trace_pupil(packed, Px, Py, fxy, out=out, spot=sums, spot_shift=shift, group_k=k_own, group_resid=gres, **kw)
for fn in (plain, verified):
    for _ in range(2): fn()
torch.cuda.synchronize()
torch.cuda.profiler.start()
plain(); torch.cuda.synchronize()
verified(); torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("prof_fx_modes ok:", 3 * npf, "rays")
