"""GPU: the reference's random pupil distribution generated on the device (olrt_pupil_random_f64) against NumPy:
RandomDistribution(seed).generate_points(n) = default_rng(seed) (PCG64) uniform draws -> sqrt(r) cos / sin(theta)
(optiland/distribution.py:132-158). The DRAWS must be NumPy's bit for bit (128-bit LCG jump-ahead + XSL-RR output);
Px, Py within 2 ulp (cos / sin come from CUDA's libdevice instead of libm)."""
import numpy as np
import pytest
import torch

from optiland_pr_b200 import backend as be
from optiland_pr_b200 import samples
from optiland_pr_b200.optic import random_pupil_device

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("seed,n", [(0, 1), (1, 31), (7, 257), (12345, 100_003), (2, 3_000_001)])
def test_device_random_pupil_matches_numpy(seed, n):
    Px, Py, raw = random_pupil_device(n, seed, "cuda", want_raw=True)
    rng = np.random.default_rng(seed)
    r = rng.uniform(size=n)
    th = rng.uniform(0, 2 * np.pi, size=n)
    raw = raw.cpu().numpy()
    assert np.array_equal(raw[0], r) and np.array_equal(raw[1], th)       # the generator's stream, bit for bit
    x, y = np.sqrt(r) * np.cos(th), np.sqrt(r) * np.sin(th)
    ulp = np.spacing(np.maximum(np.abs(x), np.abs(y)))
    assert np.max(np.abs(Px.cpu().numpy() - x) / ulp) <= 2.0
    assert np.max(np.abs(Py.cpu().numpy() - y) / ulp) <= 2.0


def test_optic_trace_random_distribution_uses_the_device_generator():
    """Optic.trace(Hx, Hy, wl, num_rays, "random", seed): the reference's call shape; same spot as the host grid."""
    be.set_precision("float64")
    from optiland_pr_b200 import spot
    from optiland_pr_b200.optic import pupil_points
    lens = samples.EvenAsphereDoubleGauss()
    lens.surface_group.record = "none"
    n, seed = 200_000, 5
    rays = lens.trace([0.0], [1.0], 0.5876, n, "random", seed=seed, spot=True)
    got = spot.spot_stats(rays.spot_sums)["rms"].cpu().numpy()
    px, py = pupil_points("random", n, seed)
    ref = lens.trace([0.0], [1.0], 0.5876, distribution=(px, py), spot=True)
    want = spot.spot_stats(ref.spot_sums)["rms"].cpu().numpy()
    assert np.allclose(got, want, rtol=1e-12, atol=0)
    assert np.max(np.abs(rays.x.cpu().numpy() - ref.x.cpu().numpy())) <= 1e-12
