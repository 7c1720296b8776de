import sys; sys.path.insert(0, "/root/repo")
import numpy as np, torch, cProfile, pstats, time
from optiland_pr_b200 import backend as be, samples, spot
be.set_device("cuda:0"); be.set_precision("float64")
lens = samples.EvenAsphereDoubleGauss(); lens.surface_group.record = "none"
n = 1 << 24
Hx, Hy = [0.0, 0.0, 0.0], [0.0, 10 / 14, 1.0]
host = torch.empty(3, dtype=torch.float64).pin_memory()
def full():
    rr = lens.trace(Hx, Hy, 0.5876, n, "random", seed=1, spot=True)
    host.copy_(spot.spot_stats(rr.spot_sums_host)["rms"])
for _ in range(5): full()
pr = cProfile.Profile(); pr.enable()
for _ in range(20): full()
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(22)
