import cProfile, pstats, time, numpy as np, torch
from optiland_pr_b200 import samples, spot
dev = torch.device("cuda", 0)
lens = samples.CookeTriplet(); lens.surface_group.record = "none"
n1 = 1_000_000
rng = np.random.default_rng(0)
r_, th = rng.uniform(size=n1), rng.uniform(0, 2*np.pi, size=n1)
Px = torch.from_numpy(np.sqrt(r_)*np.cos(th)).to(dev); Py = torch.from_numpy(np.sqrt(r_)*np.sin(th)).to(dev)
fields = lens.get_field_coords(); Hx=[f[0] for f in fields]; Hy=[f[1] for f in fields]
def c1():
    rms=[]
    for w in lens.wavelengths:
        rr = lens.trace(Hx, Hy, w, distribution=(Px, Py), spot=True)
        rms.append(spot.spot_stats(rr.spot_sums)["rms"])
    return torch.stack(rms).cpu()
for _ in range(5): c1()
torch.cuda.synchronize(); t=time.perf_counter()
for _ in range(20): c1()
torch.cuda.synchronize(); print("ms per c1", (time.perf_counter()-t)/20*1e3)
pr = cProfile.Profile(); pr.enable()
for _ in range(20): c1()
pr.disable(); pstats.Stats(pr).sort_stats("cumulative").print_stats(45)
