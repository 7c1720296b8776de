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


@pytest.mark.parametrize("rings", [0, 1, 6, 37, 600])
def test_hexapolar_pupil_on_the_device_matches_numpy(rings):
    """olrt_pupil_hexapolar_f64 against HexagonalDistribution.generate_points as the reference writes it with the NumPy
    backend (optiland/distribution.py:200-220): same point count and order; radii and angles are linspace values (bit
    for bit, checked through r = hypot and the ring structure), Px / Py within 2 ulp of r cos / r sin from libm."""
    from optiland_pr_b200.optic import hexapolar_pupil_device
    x = np.zeros([1])
    y = np.zeros([1])
    r = np.linspace(0, 1, rings + 1)
    for i in range(rings):
        num_theta = 6 * (i + 1)
        theta = np.linspace(0, 2 * np.pi, num_theta + 1)[:-1]
        x = np.concatenate([x, r[i + 1] * np.cos(theta)])
        y = np.concatenate([y, r[i + 1] * np.sin(theta)])
    Px, Py = hexapolar_pupil_device(rings, "cuda:0")
    gx, gy = Px.cpu().numpy(), Py.cpu().numpy()
    assert gx.shape == x.shape == (1 + 3 * rings * (rings + 1),)
    # 2 ulp of the radius (cos / sin are within an ulp or two of libm's; near their zeros that is absolute, not relative)
    tol = 2.0 * np.spacing(np.maximum(np.hypot(x, y), 1e-300))
    assert np.all(np.abs(gx - x) <= tol) and np.all(np.abs(gy - y) <= tol)
    if rings:
        assert gx[0] == 0.0 and gy[0] == 0.0 and gx[1] == r[1] and gy[1] == 0.0      # theta = 0: exactly (r, 0)
        assert gx[-6 * rings] == 1.0                                                  # the last ring has radius exactly 1


def test_optic_trace_with_a_large_hexapolar_grid_uses_the_device_generator():
    """Optic.trace(..., num_rays=rings, "hexapolar") with >= 4096 points: same rays as with the host-built grid."""
    from optiland_pr_b200 import backend as be, samples
    from optiland_pr_b200.optic import pupil_points
    be.set_device("cuda:0")
    be.set_precision("float64")
    lens = samples.CookeTriplet()
    lens.surface_group.record = "none"
    a = lens.trace([0.0], [0.7], 0.55, 40, "hexapolar")                  # 4921 points: device generator
    px, py = pupil_points("hexapolar", 40)
    b = lens.trace([0.0], [0.7], 0.55, distribution=(px, py))            # the host's NumPy grid
    assert a.x.numel() == b.x.numel() == 4921
    assert float((a.x - b.x).abs().max()) <= 1e-12 and float((a.opd - b.opd).abs().max()) <= 1e-11


@pytest.mark.parametrize("n", [0, 1, 2, 7, 64, 1001, 2049])
def test_uniform_pupil_on_the_device_matches_numpy_bit_for_bit(n):
    """olrt_pupil_uniform_f64 against UniformDistribution.generate_points as the reference writes it with the NumPy
    backend (optiland/distribution.py:174-185): the same points, in the same order, bit for bit (mask included)."""
    from optiland_pr_b200.optic import uniform_pupil_device
    x = np.linspace(-1, 1, n)
    x, y = np.meshgrid(x, x)
    r2 = x**2 + y**2
    wx, wy = x[r2 <= 1], y[r2 <= 1]
    Px, Py = uniform_pupil_device(n, "cuda:0")
    assert np.array_equal(Px.cpu().numpy(), wx) and np.array_equal(Py.cpu().numpy(), wy)


def test_optic_trace_with_a_large_uniform_grid_uses_the_device_generator():
    """Optic.trace(..., num_rays=n, "uniform") with n*n >= 4096: the same rays as with the host-built grid, bit for
    bit (the pupil points are identical, so the launch is the same launch)."""
    from optiland_pr_b200 import backend as be, samples
    from optiland_pr_b200.optic import pupil_points
    be.set_device("cuda:0")
    be.set_precision("float64")
    lens = samples.CookeTriplet()
    lens.surface_group.record = "none"
    a = lens.trace([0.0], [0.7], 0.55, 80, "uniform")                    # 80 x 80 grid: device generator
    px, py = pupil_points("uniform", 80)
    b = lens.trace([0.0], [0.7], 0.55, distribution=(px, py))            # the host's NumPy grid
    assert a.x.numel() == b.x.numel() == px.size
    assert bool((a.x == b.x).all()) and bool((a.opd == b.opd).all())
