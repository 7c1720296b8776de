import os, sys, time, cProfile, pstats
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import optiland_pr_b200 as ol
from optiland_pr_b200 import samples
from optiland_pr_b200.autograd import _rms_merit
ol.backend.set_device("cuda:0"); ol.backend.set_precision("float64")
dev = torch.device("cuda")
lens4 = samples.CookeTriplet(); lens4.surface_group.record = "none"
sfs = lens4.surface_group.surfaces; pars = []
for k in range(1, 7):
    p_ = torch.nn.Parameter(torch.tensor(float(sfs[k].geometry.radius), dtype=torch.float64, device=dev)); lens4.set_radius(p_, k); pars.append(p_)
opt = torch.optim.Adam(pars, lr=1e-6)
n4 = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
rng = np.random.default_rng(0)
r_, th = rng.uniform(size=n4), rng.uniform(0, 2 * np.pi, size=n4)
Px4 = torch.from_numpy(np.sqrt(r_) * np.cos(th)).to(dev); Py4 = torch.from_numpy(np.sqrt(r_) * np.sin(th)).to(dev)
fields4 = lens4.get_field_coords(); wls4 = list(lens4.wavelengths)
def c4():
    opt.zero_grad()
    loss = sum(_rms_merit(lens4, fields4, w_, Px4, Py4) for w_ in wls4)
    loss.backward(); opt.step(); return loss
for _ in range(5): c4()
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(10): c4()
torch.cuda.synchronize(); print("ms/step", (time.perf_counter() - t0) * 100)
pr = cProfile.Profile(); pr.enable()
for _ in range(10): c4()
torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats('tottime').print_stats(22)
