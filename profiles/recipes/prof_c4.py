"""One C4 optimisation step (Cooke, 9 pairs x 1 M rays, Adam) under torch.profiler: kernel list and GPU busy time."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import numpy as np, torch
from optiland_pr_b200 import backend as be, samples
from optiland_pr_b200.autograd import rms_merit_sharded
from optiland_pr_b200.optic import random_pupil_device
be.set_device("cuda:0"); be.set_precision("float64")
dev = torch.device("cuda:0")
lens = samples.CookeTriplet(); lens.surface_group.record = "none"
sfs = lens.surface_group.surfaces
pars = []
for k in range(1, 7):
    p_ = torch.nn.Parameter(torch.tensor(float(sfs[k].geometry.radius), dtype=torch.float64, device=dev)); lens.set_radius(p_, k); pars.append(p_)
for k in range(1, 7):
    p_ = torch.nn.Parameter(torch.tensor(0.0, dtype=torch.float64, device=dev)); lens.set_conic(p_, k); pars.append(p_)
opt = torch.optim.Adam(pars, lr=1e-6, fused=("fused" in sys.argv))
Px, Py = random_pupil_device(1_000_000, 0, dev)
fields, wls = lens.get_field_coords(), list(lens.wavelengths)
def step():
    opt.zero_grad()
    loss = rms_merit_sharded(lens, fields, wls, Px, Py, params=pars)
    opt.step()
    return loss
for _ in range(5): step()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(20): step()
torch.cuda.synchronize()
print("step %.3f ms wall" % ((time.perf_counter() - t0) * 1e3 / 20))
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for _ in range(5): step()
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=25, max_name_column_width=70))
import cProfile, pstats
pr = cProfile.Profile(); pr.enable()
for _ in range(20): step()
torch.cuda.synchronize()
pr.disable()
st = pstats.Stats(pr)
st.sort_stats("cumulative").print_stats(12)
st.print_callers("item")
st.print_callers("tolist")
