import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
from test_gpu_fuzz import _wild_rays
from oracle import trace_np
from optiland_pr_b200 import samples
from optiland_pr_b200.apertures import RadialAperture
from optiland_pr_b200.prescription import pack_surface_group
from optiland_pr_b200.trace import fused_trace
lens = samples.EvenAsphereDoubleGauss()
lens.surface_group.surfaces[2].aperture = RadialAperture(r_max=7.5, r_min=0.25)
wl = lens.wavelengths
packed = pack_surface_group(lens.surface_group.surfaces, wl)
n = 100000
rays_in = _wild_rays(n, 11, wl)
ora = trace_np.trace_packed(packed, trace_np.Rays(*rays_in))
ref = np.array(ora["records"])
res = fused_trace(packed, [torch.tensor(rays_in[q], device="cuda") for q in range(8)])
rec = res.record.cpu().numpy()
print("nr", res.nr_iters, ora["nr_iters"], "fixups", res.fixups)
for s in range(ref.shape[0]):
    mm = np.isnan(rec[s, 0]) != np.isnan(ref[s, 0])
    if mm.any():
        k = np.nonzero(mm)[0]
        print("surface", s, "mismatches", k.size, "first", k[:3])
        for j in k[:3]:
            print("   gpu", rec[s, :, j]); print("   ref", ref[s, :, j])
            print("   prev gpu", rec[s-1, :, j]); print("   prev ref", ref[s-1, :, j])
