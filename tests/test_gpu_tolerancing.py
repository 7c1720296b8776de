"""GPU: Monte-Carlo tolerancing through ONE fused launch (olrt_trace_pupil_* with a prescription batch) against the
NumPy oracle traced trial by trial, the way the reference's MonteCarlo.run does (tolerancing/monte_carlo.py:60-123)."""
import numpy as np
import pytest
import torch

from oracle import raygen_np, trace_np
from optiland_pr_b200 import backend as be
from optiland_pr_b200 import samples
from optiland_pr_b200.prescription import pack_surface_group
from optiland_pr_b200.tolerancing import DistributionSampler, MonteCarloSpot, Perturbation

pytestmark = pytest.mark.gpu
HX, HY = [0.0, 0.0], [0.7, 1.0]
WL = 0.5876


def _mc(lens):
    s = lens.surface_group.surfaces
    perts = [Perturbation(lens, "radius", DistributionSampler("normal", seed=100 + k, loc=float(s[k].geometry.radius),
                                                              scale=1e-3 * abs(float(s[k].geometry.radius))), k)
             for k in (1, 2, 3, 5, 7, 9, 10, 11)]
    perts += [Perturbation(lens, "thickness", DistributionSampler("normal", seed=200 + k, loc=float(s[k].thickness),
                                                                  scale=0.02), k) for k in (1, 2, 3, 5, 8)]
    perts += [Perturbation(lens, "decenter", DistributionSampler("normal", seed=300 + k, loc=0.0, scale=0.1), k, axis=ax)
              for k, ax in ((1, "x"), (5, "y"), (7, "x"), (10, "y"))]
    return MonteCarloSpot(lens, perts)


def test_fused_monte_carlo_equals_oracle_per_trial():
    be.set_precision("float64")
    lens = samples.EvenAsphereDoubleGauss()
    mc = _mc(lens)
    px, py = raygen_np.random_pupil(3000, seed=8)
    rms = mc.run(6, HX, HY, WL, px, py).cpu().numpy()
    assert mc.last_path == "fused" and rms.shape == (6, 2)
    packed = pack_surface_group(lens.surface_group.surfaces, [WL], record="none")
    batch = mc.prescriptions(packed, mc.values)
    packed_o = pack_surface_group(lens.surface_group.surfaces, [WL])      # same rows, every surface recorded
    assert packed_o.presc.shape == packed.presc.shape
    surf_z = [float(s.geometry.cs.z) for s in lens.surface_group.surfaces[1:]]
    for t in range(6):
        for f in range(2):
            rays = raygen_np.generate_rays(HX[f], HY[f], px, py, WL, EPL=lens._EPL, EPD=lens._EPD,
                                           max_field=lens.max_field, surface_z=surf_z)
            packed_o.presc = batch[t:t + 1]
            with np.errstate(all="ignore"):
                o = trace_np.trace_packed(packed_o, trace_np.Rays(*rays))["records"][-1]
            x, y = np.asarray(o[0]), np.asarray(o[1])
            want = np.sqrt(np.mean((x - x.mean()) ** 2 + (y - y.mean()) ** 2))
            assert abs(rms[t, f] - want) <= 1e-9 * want, (t, f, rms[t, f], want)
    assert np.ptp(rms[:, 1]) > 1e-4        # the perturbations really change the spot


def test_chunked_trials_equal_one_launch():
    be.set_precision("float64")
    px, py = raygen_np.random_pupil(2048, seed=2)
    a = _mc(samples.EvenAsphereDoubleGauss()).run(9, HX, HY, WL, px, py).cpu().numpy()
    b = _mc(samples.EvenAsphereDoubleGauss()).run(9, HX, HY, WL, px, py, trials_per_launch=4).cpu().numpy()
    assert np.allclose(a, b, rtol=1e-12, atol=0)
