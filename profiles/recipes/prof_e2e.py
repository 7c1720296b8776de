"""Where the e2e step (Optic.trace with a named distribution) spends its time. usage: python profiles/recipes/prof_e2e.py"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import numpy as np, torch
from optiland_pr_b200 import backend as be, samples, spot
from optiland_pr_b200.optic import random_pupil_device
be.set_device("cuda:0"); be.set_precision("float64")
dev = torch.device("cuda:0")
lens = samples.EvenAsphereDoubleGauss(); lens.surface_group.record = "none"
n = 1 << 24
Hx, Hy = [0.0, 0.0, 0.0], [0.0, 10 / 14, 1.0]

def tm(fn, k=8):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record()
    for _ in range(k): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / k, (time.perf_counter() - t0) * 1e3 / k

print("pupil generation (device, 16.7 M points): %.3f ms (wall %.3f)" % tm(lambda: random_pupil_device(n, 1, dev)))
Px, Py = random_pupil_device(n, 1, dev)
print("Optic.trace(distribution=(device Px, Py), spot=True): %.3f ms (wall %.3f)" %
      tm(lambda: lens.trace(Hx, Hy, 0.5876, distribution=(Px, Py), spot=True)))
host = torch.empty(3, dtype=torch.float64).pin_memory()
def full():
    rr = lens.trace(Hx, Hy, 0.5876, n, "random", seed=1, spot=True)
    host.copy_(spot.spot_stats(rr.spot_sums_host)["rms"])
print("full e2e step: %.3f ms (wall %.3f)" % tm(full))
import cProfile, pstats
pr = cProfile.Profile(); pr.enable()
for _ in range(5): full()
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(25)

# ---- what the launch inside Optic.trace costs, piece by piece
from optiland_pr_b200 import _lib as L
from optiland_pr_b200.trace import fused_trace_pupil
nf = 3
packed = lens.surface_group.pack([0.5876])
fxy_np, z_start = lens._field_offsets(np.asarray(Hx), np.asarray(Hy))
fxy = torch.from_numpy(fxy_np).to(dev)
kw = dict(EPD=lens._EPD, EPL=lens._EPL, z_start=z_start, wavelength=0.5876, dtype=torch.float64)
out = torch.empty((8, nf * n), dtype=torch.float64, device=dev)
dir0 = torch.empty((3, nf * n), dtype=torch.float64, device=dev)
sums = torch.zeros((nf, L.SPOT_SUMS), dtype=torch.float64, device=dev)
shift = torch.zeros((nf, 2), dtype=torch.float64, device=dev)
resid = torch.zeros((packed.n_surfaces, L.RESID_STRIDE), dtype=torch.int64, device=dev)
st = torch.zeros(1, dtype=torch.int32, device=dev)
for label, a in (("out", dict(out=out)), ("out+spot", dict(out=out, spot=sums, spot_shift=shift)),
                 ("out+spot+resid", dict(out=out, spot=sums, spot_shift=shift, resid_out=resid, status_t=st)),
                 ("out+dir0+spot+resid", dict(out=out, dir0=dir0, spot=sums, spot_shift=shift, resid_out=resid, status_t=st)),
                 ("spot only", dict(spot=sums, spot_shift=shift))):
    print("%-22s %.3f ms" % (label, tm(lambda: fused_trace_pupil(packed, Px, Py, fxy, **a, **kw))[0]))
print("alloc out+dir0: %.3f ms" % tm(lambda: (torch.empty((8, nf * n), dtype=torch.float64, device=dev), torch.empty((3, nf * n), dtype=torch.float64, device=dev)))[0])
