"""Small end-to-end case for compute-sanitizer (memcheck / racecheck): every kernel family once, tiny sizes.
  - trace_fx (array mode, dead rays -> the strict IEEE fix-up pass), trace_fk (strict), interpreter (freeform plate)
  - pupil mode with fused spot sums, group mode (Monte-Carlo batch), STREAMED pupil grid (ready words)
  - forward with tile-blocked checkpoints + reduced-algebra adjoint, general adjoint on records
  - spot sums, OPD epilogue, device random pupil
usage: compute-sanitizer --tool memcheck python profiles/recipes/sanitize_case.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
from optiland_pr_b200 import backend as be, samples, spot
from optiland_pr_b200.autograd import trace_differentiable
from optiland_pr_b200.optic import random_pupil_device
from optiland_pr_b200.prescription import pack_surface_group
from optiland_pr_b200.tolerancing import DistributionSampler, MonteCarloSpot, Perturbation
from optiland_pr_b200.trace import fused_trace
from helpers import load_golden, mirror_packed

be.set_device("cuda:0"); be.set_precision("float64")
dev = torch.device("cuda:0")
done = []
# 1. array mode on golden rays: fx (tolerance), fk (strict), interpreter, fp32
for name in ("cooke_triplet", "even_asphere_double_gauss", "freeform_plate_fringe", "three_mirror"):
    g = load_golden(name)
    _, packed = mirror_packed(name, g["wavelengths"])
    rays = [torch.tensor(g["rays_in"][q], dtype=torch.float64, device=dev) for q in range(8)]
    for fast in (True, False):
        fused_trace(packed, rays, fast=fast)
    fused_trace(packed, [t.float() for t in rays])
    done.append(name)
# 2. wild rays (misses, TIR, clipped): the fix-up list is long
lens = samples.EvenAsphereDoubleGauss()
rng = np.random.default_rng(1)
n = 20_000
x, y = rng.uniform(-14, 14, n), rng.uniform(-14, 14, n)
Lc, Mc = rng.uniform(-0.5, 0.5, n), rng.uniform(-0.5, 0.5, n)
wild = np.stack([x, y, np.full(n, -8.0), Lc, Mc, np.sqrt(1 - Lc * Lc - Mc * Mc), np.ones(n), np.full(n, 0.5876)])
packed = pack_surface_group(lens.surface_group.surfaces, [0.5876])
r = fused_trace(packed, [torch.tensor(wild[q], device=dev) for q in range(8)], fast=True)
done.append(f"wild rays, {r.fixups} re-traced")
# 3. pupil mode: resident, chunked, STREAMED host grid with spot sums; Monte-Carlo group mode
lens.surface_group.record = "none"
px, py = random_pupil_device(70_001, 3, dev)
rr = lens.trace([0.0, 0.0], [0.0, 1.0], 0.5876, distribution=(px, py), spot=True)
lens.pipeline_min_rays = 1 << 14
hx, hy = px.cpu().pin_memory(), py.cpu().pin_memory()
rs = lens.trace([0.0, 0.0], [0.0, 1.0], 0.5876, distribution=(hx, hy), spot=True)
assert torch.equal(rr.x, rs.x)
done.append("pupil mode resident + streamed")
s = lens.surface_group.surfaces
mc = MonteCarloSpot(lens, [Perturbation(lens, "radius", DistributionSampler("normal", seed=1, loc=float(s[1].geometry.radius), scale=0.05), 1),
                           Perturbation(lens, "decenter", DistributionSampler("normal", seed=2, loc=0.0, scale=0.1), 5, axis="y")])
mc.run(5, [0.0, 0.0], [0.7, 1.0], 0.5876, px[:4000], py[:4000])
sd = spot.SpotDiagram(lens, num_rings=4)
done.append("group mode (Monte Carlo, SpotDiagram)")
# 4. differentiable traces: blocked checkpoints + reduced adjoint, records + general adjoint
rays = lens.generate_rays(torch.zeros(9000, dtype=torch.float64, device=dev), torch.full((9000,), 0.7, dtype=torch.float64, device=dev),
                          px[:9000], py[:9000], 0.5876)
ins = rays.components(); ins[7] = None
for record, blocked, generic in (("none", True, False), ("all", False, False), ("all", False, True)):
    pk = pack_surface_group(lens.surface_group.surfaces, [0.5876], record=record)
    presc = torch.tensor(pk.presc[0], dtype=torch.float64, device=dev, requires_grad=True)
    leaf = [t.detach().clone().requires_grad_(q < 6) if t is not None else None for q, t in enumerate(ins)]
    out, rec, _ = trace_differentiable(pk, presc, leaf, fast=True, blocked=blocked, generic_adjoint=generic)
    (out[0] ** 2 + out[1] ** 2).sum().backward()
done.append("forward + backward")
# 5. wavefront epilogue
from optiland_pr_b200.wavefront import Wavefront
ck = samples.CookeTriplet()
g = load_golden("wavefront_cooke")
ck.set_paraxial(EPL=float(g["EPL"]), EPD=float(g["EPD"]), XPL=float(g["XPL"]))
Wavefront(ck, fields=[(0.0, 1.0)], wavelengths=[0.55], num_rays=8).get_data((0.0, 1.0), 0.55)
done.append("OPD epilogue")
torch.cuda.synchronize()
print("sanitize_case ok:", "; ".join(done))
