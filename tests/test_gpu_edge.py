"""GPU edge cases (rim rays, vertex hits, a == 0, N ~ 0, empty and ragged batches) against the NumPy
oracle, which tests/test_oracle_vs_reference_live.py pins to the unmodified reference on the same constructions.
Strict mode: bit for bit. Tolerance mode: <= 1e-10 with identical masks and NaN patterns."""
import numpy as np
import pytest
import torch

from oracle import trace_np
from optiland_pr_b200 import backend as be
from optiland_pr_b200.apertures import RadialAperture
from optiland_pr_b200.optic import Optic
from optiland_pr_b200.prescription import pack_surface_group
from optiland_pr_b200.trace import fused_trace

pytestmark = pytest.mark.gpu
INF = float("inf")


def _singlet(**kw):
    be.set_precision("float64")
    lens = Optic("edge")
    lens.add_surface(index=0, radius=INF, thickness=INF)
    lens.add_surface(index=1, thickness=5.0, material="N-SF11", is_stop=True, **kw)
    lens.add_surface(index=2, radius=INF, thickness=20.0)
    lens.add_surface(index=3)
    lens.add_wavelength(0.55, is_primary=True)
    return lens


def _check(lens, rays_in, fast, keep=None):
    packed = pack_surface_group(lens.surface_group.surfaces, [0.55])
    with np.errstate(all="ignore"):
        want = np.array(trace_np.trace_packed(packed, trace_np.Rays(*rays_in))["records"])
    res = fused_trace(packed, [torch.tensor(rays_in[q], device="cuda") for q in range(8)], fast=fast)
    got = res.record.cpu().numpy()
    if keep is not None:
        got, want = got[:, :, keep], want[:, :, keep]
    assert np.array_equal(np.isnan(got), np.isnan(want))
    assert np.array_equal(got[:, 6] == 0, want[:, 6] == 0)
    if fast:
        m = np.isfinite(want)
        assert np.array_equal(got[~m], want[~m], equal_nan=True)      # NaN and +-inf in the same places
        assert np.max(np.abs(got[m] - want[m]) / np.maximum(np.abs(want[m]), 1.0), initial=0.0) <= 1e-10
    else:
        for q in (0, 1, 2, 3, 4, 5, 7):
            assert np.array_equal(got[:, q], want[:, q], equal_nan=True), q
        assert np.allclose(got[:, 6], want[:, 6], rtol=1e-12, atol=0, equal_nan=True)
    return want


@pytest.mark.parametrize("fast", [False, True], ids=["strict", "tolerance"])
def test_paraboloid_axis_rays_a_equals_zero(fast):
    """k = -1, rays along the axis: a = k N^2 + L^2 + M^2 + N^2 == 0 -> t = -c / b (standard.py:136-138)."""
    lens = _singlet(radius=30.0, conic=-1.0)
    n = 64
    rng = np.random.default_rng(5)
    x, y = rng.uniform(-4, 4, n), rng.uniform(-4, 4, n)
    x[0] = y[0] = 0.0
    rays_in = np.stack([x, y, np.full(n, -3.0), np.zeros(n), np.zeros(n), np.ones(n), np.ones(n), np.full(n, 0.55)])
    want = _check(lens, rays_in, fast)
    assert np.isfinite(want[0, 2]).all()


@pytest.mark.parametrize("fast", [False, True], ids=["strict", "tolerance"])
def test_flat_standard_surface_grazing_rays_clamp(fast):
    """StandardGeometry with an infinite radius: t = -z / where(|N| > 1e-14, N, 1e-14) (standard.py:100-103)."""
    lens = _singlet(radius=INF)
    n = 32
    N = np.concatenate([np.array([0.0, 1e-15, -1e-15, 1e-14, 2e-14, 1e-9]), np.linspace(0.01, 1.0, n - 6)])
    L = np.sqrt(np.maximum(0.0, 1.0 - N * N))
    rays_in = np.stack([np.zeros(n), np.linspace(-1, 1, n), np.full(n, -1.0), L, np.zeros(n), N, np.ones(n),
                        np.full(n, 0.55)])
    # The six grazing rays leave the second plane exactly AT the critical angle (1 - u^2 (1 - dot^2) cancels to 0 or
    # to a few 1e-17): whether N' is 0 (t = inf at the image) or 3e-9 is decided by the last rounding. In tolerance
    # mode that radicand is inside the guard band, so those rays are re-traced by the strict IEEE kernel and carry the
    # reference's decision like every other ray.
    _check(lens, rays_in, fast)


@pytest.mark.parametrize("fast", [False, True], ids=["strict", "tolerance"])
def test_rim_rays_on_the_aperture_edge(fast):
    lens = _singlet(radius=40.0, aperture=RadialAperture(r_max=3.0, r_min=0.5))
    r = np.array([0.0, 0.5, np.nextafter(0.5, 0), np.nextafter(0.5, 1), 3.0, np.nextafter(3.0, 0), np.nextafter(3.0, 4),
                  1.0, 2.999999999, 3.000000001])
    n = r.size
    rays_in = np.stack([r, np.zeros(n), np.full(n, -2.0), np.zeros(n), np.zeros(n), np.ones(n), np.ones(n),
                        np.full(n, 0.55)])
    want = _check(lens, rays_in, fast)
    assert (want[0, 6] == 0).any() and (want[0, 6] != 0).any()


@pytest.mark.parametrize("fast", [False, True], ids=["strict", "tolerance"])
@pytest.mark.parametrize("n", [0, 1, 31, 257])
def test_empty_and_ragged_batches(n, fast):
    lens = _singlet(radius=25.0)
    rng = np.random.default_rng(n)
    x, y = rng.uniform(-3, 3, n), rng.uniform(-3, 3, n)
    rays_in = np.stack([x, y, np.full(n, -2.0), np.zeros(n), np.zeros(n), np.ones(n), np.ones(n), np.full(n, 0.55)])
    packed = pack_surface_group(lens.surface_group.surfaces, [0.55])
    res = fused_trace(packed, [torch.tensor(rays_in[q], device="cuda") for q in range(8)], fast=fast)
    assert tuple(res.out.shape) == (8, n)
    if n:
        _check(lens, rays_in, fast)


@pytest.mark.parametrize("fast", [True, False])
@pytest.mark.parametrize("kind", ["zernike_m0", "zernike_tilt", "odd_asphere"])
def test_rays_on_the_axis_of_zernike_and_odd_asphere_surfaces(kind, fast):
    """The interpreter's unguarded arithmetic (KIND_GFAST) on surfaces whose evaluation takes a square root of EXACTLY
    zero on the axis (rho of a Zernike term, r of an odd asphere): an unguarded sqrt(0) is a NaN that the reference's
    own `where(|f'| > 1e-14, ...)` select could swallow -- a rotationally symmetric Zernike surface (m = 0 terms only)
    has a finite sag there, so the ray would go on with a wrong finite intersection. Those two roots keep their guard;
    here rays through the vertex, 1e-300 and 1e-9 off it, and ordinary ones, tilted frame included.
    Zernike terms go through cos / sin(m phi) (libm in the oracle, angle addition in the kernel): <= 1e-12 even in strict
    mode, as for every trigonometric surface."""
    if kind == "odd_asphere":
        # (no C_1 r term: with one the vertex is a cone tip, the slope C_1 x / r is discontinuous there and the last
        # ulp of x decides the reference's own answer -- nothing a tolerance can be stated for)
        kw = dict(radius=40.0, conic=-0.2, surface_type="odd_asphere", coefficients=[0.0, 2.0e-4, -3.0e-5], ry=0.01)
    elif kind == "zernike_m0":      # fringe 1, 4, 9: piston, defocus, spherical -- no azimuthal dependence
        kw = dict(radius=50.0, conic=0.1, surface_type="zernike", zernike_type="fringe", norm_radius=8.0,
                  coefficients=[1.0e-3, 0.0, 0.0, 2.0e-2, 0.0, 0.0, 0.0, 0.0, -5.0e-3])
    else:                           # tilt + coma terms as well, on a tilted frame
        kw = dict(radius=50.0, conic=0.1, surface_type="zernike", zernike_type="noll", norm_radius=8.0, rx=0.02,
                  coefficients=[0.0, 1.0e-2, -2.0e-2, 1.0e-2, 0.0, 3.0e-3, 4.0e-3, -2.0e-3])
    lens = _singlet(**kw)
    # (the last ray sits 1e-100 off the axis: r^2 = 2e-200 is still a normal number. A ray 1e-160 off it -- r^2 a
    # denormal -- is outside the stated domain of the tolerance mode, 2^+-400, and differs there by 4e-6 in the OPD)
    xs = np.array([0.0, 0.0, 1e-300, 1e-9, -1e-9, 0.5, -2.0, 3.0, 0.0, 1e-100])
    ys = np.array([0.0, 1e-300, 0.0, 1e-9, 2e-9, -1.5, 0.25, 1.0, 2.5, -1e-100])
    n = xs.size
    L_ = np.array([0.0, 0.0, 0.0, 0.0, 1e-3, 0.01, -0.02, 0.0, 0.0, 0.0])
    M_ = np.zeros(n)
    rays_in = [xs, ys, np.full(n, -5.0), L_, M_, np.sqrt(1 - L_ * L_ - M_ * M_), np.ones(n), np.full(n, 0.55)]
    if kind.startswith("zernike") and not fast:
        packed = pack_surface_group(lens.surface_group.surfaces, [0.55])
        with np.errstate(all="ignore"):
            want = np.array(trace_np.trace_packed(packed, trace_np.Rays(*rays_in))["records"])
        got = fused_trace(packed, [torch.tensor(rays_in[q], device="cuda") for q in range(8)], fast=False).record.cpu().numpy()
        assert np.array_equal(np.isnan(got), np.isnan(want)) and np.isfinite(want).all()
        assert np.max(np.abs(got - want) / np.maximum(np.abs(want), 1.0)) <= 1e-12
    else:
        _check(lens, rays_in, fast)
