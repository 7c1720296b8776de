"""GPU: the fall-backs of the batched paths.
 * a ray that misses an asphere makes the reference run max_iter Newton updates for its whole trace
   (newton_raphson.py:137-149): more than the fused (trial x field) launch holds -> SpotDiagram / MonteCarloSpot must
   trace those pair by pair / trial by trial, with the oracle's numbers;
 * lenses or perturbations outside the fused launch (freeforms, tilts) with a pupil grid whose size is not a multiple
   of 32 (every hexapolar grid is odd) and several trials;
 * tilt perturbations."""
import numpy as np
import pytest
import torch

from oracle import raygen_np, trace_np
from optiland_pr_b200 import backend as be
from optiland_pr_b200 import samples, spot
from optiland_pr_b200.optic import pupil_points
from optiland_pr_b200.prescription import pack_surface_group
from optiland_pr_b200.tolerancing import DistributionSampler, MonteCarloSpot, Perturbation

pytestmark = pytest.mark.gpu


def _oracle_rays(lens, packed, hx, hy, px, py, wl):
    surf_z = [float(s.geometry.cs.z) for s in lens.surface_group.surfaces[1:]]
    rays = raygen_np.generate_rays(hx, hy, px, py, wl, EPL=lens._EPL, EPD=lens._EPD, max_field=lens.max_field,
                                   surface_z=surf_z)
    with np.errstate(all="ignore"):
        o = trace_np.trace_packed(packed, trace_np.Rays(*rays))["records"][-1]
    return np.asarray(o[0]), np.asarray(o[1]), np.asarray(o[6])


def test_spot_diagram_with_rays_that_miss_falls_back_to_per_pair():
    be.set_precision("float64")
    lens = samples.EvenAsphereDoubleGauss()
    rng = np.random.default_rng(11)
    r_, th = rng.uniform(size=4000), rng.uniform(0, 2 * np.pi, size=4000)
    px, py = np.sqrt(r_) * np.cos(th), np.sqrt(r_) * np.sin(th)
    px[:40] *= 6.0       # far outside the pupil: these rays miss a surface (NaN)
    py[:40] *= 6.0
    wl = 0.5876
    sd = spot.SpotDiagram(lens, fields=[(0.0, 0.0), (0.0, 1.0)], wavelengths=[wl], distribution=(px, py))
    assert sd.path == "per_pair"
    rms = np.array([[float(v) for v in row] for row in sd.rms_spot_radius()])
    packed = pack_surface_group(lens.surface_group.surfaces, [wl])
    n_nan = 0
    for f, (hx, hy) in enumerate([(0.0, 0.0), (0.0, 1.0)]):
        x, y, i = _oracle_rays(lens, packed, hx, hy, px, py, wl)
        ok = i > 0                       # spot_diagram.py:427-429 (NaN intensity compares False)
        n_nan += int(np.isnan(x).sum())
        cx, cy = x[ok].mean(), y[ok].mean()
        want = np.sqrt(np.mean((x[ok] - cx) ** 2 + (y[ok] - cy) ** 2))
        assert abs(rms[f, 0] - want) <= 1e-9 * want, (f, rms[f, 0], want)
    assert n_nan > 0


def test_monte_carlo_with_rays_that_miss_falls_back_trial_by_trial():
    be.set_precision("float64")
    lens = samples.EvenAsphereDoubleGauss()
    s = lens.surface_group.surfaces
    perts = [Perturbation(lens, "radius", DistributionSampler("normal", seed=5, loc=float(s[1].geometry.radius),
                                                              scale=0.05), 1),
             Perturbation(lens, "decenter", DistributionSampler("normal", seed=6, loc=0.0, scale=0.1), 5, axis="y")]
    mc = MonteCarloSpot(lens, perts)
    px, py = raygen_np.random_pupil(1500, seed=4)
    px, py = px.copy(), py.copy()
    px[:10] *= 6.0
    py[:10] *= 6.0
    HX, HY, wl = [0.0], [1.0], 0.5876
    rms = mc.run(3, HX, HY, wl, px, py).cpu().numpy()
    assert mc.last_path == "fused" and mc.fallback_trials >= 3
    packed = pack_surface_group(lens.surface_group.surfaces, [wl])
    batch = mc.prescriptions(packed, mc.values)
    for t in range(3):
        packed.presc = batch[t:t + 1]
        x, y, _ = _oracle_rays(lens, packed, 0.0, 1.0, px, py, wl)
        # RayOperand.rms_spot_size counts every ray (ray.py:336-340): a NaN ray makes the merit NaN
        assert np.isnan(x).any() and np.isnan(rms[t, 0])
    # the same trials without the wild rays: finite, equal to the oracle
    mc2 = MonteCarloSpot(lens, [Perturbation(lens, p.variable_type, DistributionSampler("normal", seed=sd, loc=loc,
                                                                                        scale=sc), p.surface_number,
                                             axis=p.axis)
                                for p, sd, loc, sc in zip(perts, (5, 6), (float(s[1].geometry.radius), 0.0), (0.05, 0.1))])
    rms2 = mc2.run(3, HX, HY, wl, px[10:], py[10:]).cpu().numpy()
    assert mc2.fallback_trials == 0 and np.isfinite(rms2).all()


def test_monte_carlo_on_a_freeform_lens_hexapolar_grid_and_tilts():
    """FreeformPlate (Zernike + odd asphere + tilted frames: the interpreter kernel), 127-point hexapolar grid,
    4 trials, radius + decenter + TILT perturbations: every trial equals the oracle tracing the lens packed with that
    trial's values applied to the objects."""
    be.set_precision("float64")
    lens = samples.FreeformPlate()
    s = lens.surface_group.surfaces
    perts = [Perturbation(lens, "radius", DistributionSampler("normal", seed=21, loc=float(s[3].geometry.radius),
                                                              scale=0.02), 3),
             Perturbation(lens, "decenter", DistributionSampler("normal", seed=22, loc=float(s[2].geometry.cs.x),
                                                                scale=0.05), 2, axis="x"),
             Perturbation(lens, "tilt", DistributionSampler("normal", seed=23, loc=float(s[3].geometry.cs.ry),
                                                            scale=0.003), 3, axis="y"),
             Perturbation(lens, "tilt", DistributionSampler("normal", seed=24, loc=0.0, scale=0.002), 4, axis="x")]
    mc = MonteCarloSpot(lens, perts)
    px, py = pupil_points("hexapolar", 6)
    px, py = np.asarray(px) * 0.5, np.asarray(py) * 0.5      # inside the TIR rim: finite merits
    assert px.size % 32 != 0
    HX, HY, wl = [0.0, 0.0], [0.0, 1.0], 0.55
    rms = mc.run(4, HX, HY, wl, px, py).cpu().numpy()
    assert mc.last_path == "per_trial" and rms.shape == (4, 2)
    for t in range(4):
        pl = samples.FreeformPlate()
        ps = pl.surface_group.surfaces
        ps[3].geometry.radius = float(mc.values[t, 0])
        ps[2].geometry.cs.x = float(mc.values[t, 1])
        ps[3].geometry.cs.ry = float(mc.values[t, 2])
        ps[4].geometry.cs.rx = float(mc.values[t, 3])
        packed = pack_surface_group(ps, [wl])
        for f in range(2):
            x, y, _ = _oracle_rays(pl, packed, HX[f], HY[f], px, py, wl)
            want = np.sqrt(np.mean((x - x.mean()) ** 2 + (y - y.mean()) ** 2))
            assert np.isfinite(want) and abs(rms[t, f] - want) <= 1e-9 * want, (t, f, rms[t, f], want)
    assert np.ptp(rms[:, 1]) > 1e-6
