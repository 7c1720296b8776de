import os, sys, time, cProfile, pstats
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import optiland_pr_b200 as ol
from optiland_pr_b200 import samples, spot
ol.backend.set_device("cuda:0"); ol.backend.set_precision("float64")
npf = 1 << 24
lens = samples.EvenAsphereDoubleGauss(); lens.surface_group.record = "none"; lens.pipeline_chunks = 12
rng = np.random.default_rng(1)
r_ = rng.uniform(size=npf); th = rng.uniform(0, 2*np.pi, size=npf)
Px_h = torch.from_numpy(np.sqrt(r_)*np.cos(th)).pin_memory(); Py_h = torch.from_numpy(np.sqrt(r_)*np.sin(th)).pin_memory()
Hx = [0.0, 0.0, 0.0]; Hy = [0.0, 10/14, 1.0]
rms_host = torch.empty(3, dtype=torch.float64).pin_memory()
def step():
    rr = lens.trace(Hx, Hy, 0.5876, distribution=(Px_h, Py_h), spot=True)
    rms_host.copy_(spot.spot_stats(rr.spot_sums)["rms"], non_blocking=False)
for _ in range(5): step()
torch.cuda.synchronize()
pr = cProfile.Profile(); pr.enable()
for _ in range(20): step()
pr.disable()
pstats.Stats(pr).sort_stats('tottime').print_stats(28)
