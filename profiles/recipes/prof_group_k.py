"""A/B of the Newton-count variants of the fused pupil kernel on the bench batch: batch-global counts with bit reports
(MODE 9), per-field counts with bit reports (MODE 25) given equal / the fields' own counts; fix-up list lengths.
usage: python profiles/recipes/prof_group_k.py [rays_per_field]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import numpy as np, torch
import optiland_pr_b200 as ol
from optiland_pr_b200 import _lib, samples, trace as T
from optiland_pr_b200.optic import random_pupil_device
npf = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 24
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
resid = torch.zeros((S, _lib.RESID_STRIDE), dtype=torch.int64, device=dev)
gres = torch.zeros((3, S, _lib.GROUP_RESID), dtype=torch.int64, device=dev)
k_same = torch.tensor([kmax] * 3, dtype=torch.int32, device=dev)
k_own = torch.tensor([[1, 0, 0, 0, 1, 0, 1, 0, 0, 0, 1, 0]] * 2 + [kmax], dtype=torch.int32, device=dev)

def tm(fn, k=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(k): fn()
    e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1) / k

def fix():
    return int(T._fixup_workspace(3 * npf, dev)[0].item())

cases = {
    "no verification (MODE 8)": lambda: T.fused_trace_pupil(packed, Px, Py, fxy, out=out, spot=sums, spot_shift=shift, **kw),
    "batch counts, bit reports (MODE 9)": lambda: T.fused_trace_pupil(packed, Px, Py, fxy, out=out, spot=sums, spot_shift=shift, resid_out=resid, **kw),
    "per-field counts, all = max (MODE 25)": lambda: T.fused_trace_pupil(packed, Px, Py, fxy, out=out, spot=sums, spot_shift=shift, group_k=k_same, group_resid=gres, **kw),
    "per-field counts, own (MODE 25)": lambda: T.fused_trace_pupil(packed, Px, Py, fxy, out=out, spot=sums, spot_shift=shift, group_k=k_own, group_resid=gres, **kw),
}
for name, fn in cases.items():
    ms = tm(fn)
    print(f"{name:42s} {ms:7.3f} ms   fix-up list {fix()}")
