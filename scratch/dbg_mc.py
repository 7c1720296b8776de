import sys, os
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import numpy as np, torch
from test_gpu_tolerancing import _mc, HX, HY, WL
from oracle import raygen_np, trace_np
from optiland_pr_b200 import backend as be, samples, spot, _lib as L
from optiland_pr_b200.prescription import pack_surface_group
from optiland_pr_b200.trace import fused_trace, fused_trace_pupil
be.set_precision("float64")
lens = samples.EvenAsphereDoubleGauss(); mc = _mc(lens)
px, py = raygen_np.random_pupil(3000, seed=8)
rms = mc.run(6, HX, HY, WL, px, py).cpu().numpy()
packed = pack_surface_group(lens.surface_group.surfaces, [WL], record="none")
packed_o = pack_surface_group(lens.surface_group.surfaces, [WL])
batch = mc.prescriptions(packed, mc.values)
surf_z = [float(s.geometry.cs.z) for s in lens.surface_group.surfaces[1:]]
t, f = 0, 0
rays = raygen_np.generate_rays(HX[f], HY[f], px, py, WL, EPL=lens._EPL, EPD=lens._EPD, max_field=lens.max_field, surface_z=surf_z)
packed_o.presc = batch[t:t+1]
o = trace_np.trace_packed(packed_o, trace_np.Rays(*rays))
x, y = np.asarray(o["records"][-1][0]), np.asarray(o["records"][-1][1])
rm = lambda x, y: np.sqrt(np.mean((x - x.mean())**2 + (y - y.mean())**2))
print("oracle", rm(x, y), "nr", o["nr_iters"], "nan", np.isnan(x).sum())
ins = [torch.tensor(np.asarray(rays[q]), device="cuda") for q in range(8)]
pr = torch.tensor(batch[t:t+1], device="cuda")
for fast in (False, True):
    packed.set_nr_iters([2 if lay.geom >= 2 else 0 for lay in packed.layouts])
    r = fused_trace(packed, ins, presc=pr, fast=fast)
    xo, yo = r.out[0].cpu().numpy(), r.out[1].cpu().numpy()
    print("gpu arrays fast", fast, rm(xo, yo), "nr", r.nr_iters, "maxdx", np.abs(xo - x).max())
print("mc rms", rms[t, f], "mc nr", "last packed nr?")
fxy_np, z_start = lens._field_offsets(np.asarray(HX), np.asarray(HY))
fxy = torch.tensor(fxy_np, device="cuda"); Px = torch.tensor(px, device="cuda"); Py = torch.tensor(py, device="cuda")
kw = dict(EPD=lens._EPD, EPL=lens._EPL, z_start=z_start, wavelength=WL, dtype=torch.float64)
packed.set_nr_iters([1 if lay.geom >= 2 else 0 for lay in packed.layouts])
T = 6; n = 3000
out = torch.zeros((8, T * 2 * n), dtype=torch.float64, device="cuda")
fused_trace_pupil(packed, Px, Py, fxy, out=out, presc=torch.tensor(batch, device="cuda"), **kw)
torch.cuda.synchronize()
xo = out[0, :n].cpu().numpy(); yo = out[1, :n].cpu().numpy()
print("pupil batch group0: rms", rm(xo, yo), "maxdx vs oracle", np.abs(xo - x).max(), np.abs(yo - y).max())
out1 = torch.zeros((8, 2 * n), dtype=torch.float64, device="cuda")
fused_trace_pupil(packed, Px, Py, fxy, out=out1, presc=torch.tensor(batch[:1], device="cuda"), **kw)
x1 = out1[0, :n].cpu().numpy()
print("pupil single presc group0: maxdx vs oracle", np.abs(x1 - x).max())
sums = torch.zeros((T*2, L.SPOT_SUMS), dtype=torch.float64, device="cuda")
shift = torch.zeros((T*2, 2), dtype=torch.float64, device="cuda"); shift[:, 1] = 10.0
fused_trace_pupil(packed, Px, Py, fxy, out=None, spot=sums, spot_shift=shift, presc=torch.tensor(batch, device="cuda"), **kw)
sums[:, 8:10] = shift
print("fused spot rms", spot.spot_stats(sums)["rms"][:2].cpu().numpy(), "count", sums[:2, 0].cpu().numpy())
for K in (1, 2, 3):
    packed.set_nr_iters([K if lay.geom >= 2 else 0 for lay in packed.layouts])
    fused_trace_pupil(packed, Px, Py, fxy, out=out, presc=torch.tensor(batch, device="cuda"), **kw)
    torch.cuda.synchronize()
    xo = out[0, :n].cpu().numpy(); yo = out[1, :n].cpu().numpy()
    print("K", K, "group0 rms", rm(xo, yo), "maxdx", np.abs(xo - x).max())
