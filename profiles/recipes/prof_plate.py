#!/usr/bin/env python
"""C3 plate (Zernike fringe freeform + odd asphere, tilted frames): tolerance build against the strict build on the
same rays (largest difference per component), then the time per 64 M rays like bench.py's leg.
usage: prof_plate.py [log2 rays]"""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
from optiland_pr_b200 import backend as be, samples
from optiland_pr_b200.optic import random_pupil_device
from optiland_pr_b200.trace import fused_trace

be.set_device("cuda:0"); be.set_precision("float64")
dev = torch.device("cuda:0")
for ztype in ("fringe", "noll", "standard"):
    lens = samples.FreeformPlate(ztype)
    wl = lens.primary_wavelength
    n = 1 << 21
    Px, Py = random_pupil_device(n, 2, dev)
    for f0 in lens.get_field_coords():
        rays = lens.launch_rays(np.array([f0[0]]), np.array([f0[1]]), Px, Py, wl)
        ins = rays.components(); ins[7] = None
        lens.surface_group.record = "none"
        packed = lens.surface_group.pack([float(wl)])
        a = torch.empty((8, n), dtype=torch.float64, device=dev)
        b = torch.empty((8, n), dtype=torch.float64, device=dev)
        ra = fused_trace(packed, ins, out=a, fast=True)
        rb = fused_trace(packed, ins, out=b, fast=False)
        same_nan = bool((torch.isnan(a) == torch.isnan(b)).all())
        d = (torch.nan_to_num(a) - torch.nan_to_num(b)).abs().amax(dim=1).tolist()
        print(ztype, f0, "K", ra.nr_iters, rb.nr_iters, "nan pattern equal", same_nan, "max |diff|",
              " ".join("%.1e" % v for v in d), "clip differs", int(((a[6] == 0) != (b[6] == 0)).sum()))
lg = int(sys.argv[1]) if len(sys.argv) > 1 else 26
lens = samples.FreeformPlate()
wl = lens.primary_wavelength
nz = 1 << lg
Px, Py = random_pupil_device(nz, 2, dev)
for f0 in lens.get_field_coords():
    rays = lens.launch_rays(np.array([f0[0]]), np.array([f0[1]]), Px, Py, wl)
    ins = rays.components(); ins[7] = None
    lens.surface_group.record = "none"
    packed = lens.surface_group.pack([float(wl)])
    o = torch.empty((8, nz), dtype=torch.float64, device=dev)
    fused_trace(packed, ins, out=o, fast=True)
    for fast in (True, False):
        ts = []
        for _ in range(4):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fused_trace(packed, ins, out=o, fast=fast, verify_nr=False, check_status=False)
            e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        print("field", f0, "fast" if fast else "strict", "ms per %d rays: %.2f" % (nz, min(ts[1:])))
    del o, ins, rays
