"""Decision margins of the interpreter kernel's tolerance build (csrc/trace_common.cuh: Margin): a ray
whose aperture comparison, Snell radicand, discriminant, root choice or 1e-14 switch lies within 2^-36 (relative) of
its threshold is listed for the strict build's re-trace, so its clip / TIR / miss DECISION is the reference's.
Rays are put ON the thresholds here (aperture edges to the ulp, a sweep of adjacent doubles across the critical
angle) on lenses that the reduced-algebra kernel does not take (odd asphere, tilted faces): masks and NaN patterns
must equal the strict build's -- which is bit-identical to the reference on these closed-form lenses -- ray for ray."""
import numpy as np
import pytest
import torch

from optiland_pr_b200 import backend as be
from optiland_pr_b200.apertures import (DifferenceAperture, EllipticalAperture, RadialAperture, RectangularAperture)
from optiland_pr_b200.optic import Optic
from optiland_pr_b200.prescription import pack_surface_group
from optiland_pr_b200.trace import fused_trace

pytestmark = pytest.mark.gpu
INF = float("inf")


def _singlet(first, second=None):
    be.set_precision("float64")
    lens = Optic("margins")
    lens.add_surface(index=0, radius=INF, thickness=INF)
    lens.add_surface(index=1, thickness=6.0, material="N-SF11", is_stop=True, **first)
    lens.add_surface(index=2, thickness=20.0, **(second or dict(radius=INF)))
    lens.add_surface(index=3)
    lens.add_wavelength(0.55, is_primary=True)
    return lens


def _both(lens, rays_in):
    packed = pack_surface_group(lens.surface_group.surfaces, [0.55])
    assert not packed.fastkinds()                       # the interpreter kernel, not the reduced-algebra one
    dev = [torch.tensor(np.ascontiguousarray(v), device="cuda") for v in rays_in]
    fast = fused_trace(packed, dev, fast=True)
    strict = fused_trace(packed, dev, fast=False)
    return fast, strict


def _edge_values(v):
    """v itself, its neighbours to the ulp, v (1 +- 1e-13) (inside the band), v (1 +- 1e-9) (outside)."""
    return [v, np.nextafter(v, 0), np.nextafter(v, 2 * v), v * (1 + 1e-13), v * (1 - 1e-13), v * (1 + 1e-9), v * (1 - 1e-9)]


@pytest.mark.parametrize("kind", ["radial", "rectangular", "elliptical_minus_rectangle"])
def test_rays_on_aperture_edges_take_the_strict_decision(kind):
    if kind == "radial":
        ap = RadialAperture(r_max=3.0, r_min=0.5)
        xs = _edge_values(3.0) + _edge_values(0.5) + [1.0, 2.0, 4.0]
        ys = [0.0] * len(xs)
        on_edge = 5 + 5
    elif kind == "rectangular":
        ap = RectangularAperture(-2.0, 1.5, -1.0, 0.75)
        xs = _edge_values(1.5) + [0.3] * 7 + [-v for v in _edge_values(2.0)] + [0.0, 5.0]
        ys = [0.1] * 7 + _edge_values(0.75) + [0.2] * 7 + [0.0, 0.0]
        on_edge = 5 + 5 + 5
    else:
        ap = DifferenceAperture(EllipticalAperture(4.0, 2.0), RectangularAperture(-1.0, 1.0, -0.5, 0.5))
        xs = _edge_values(4.0) + [0.0] * 7 + _edge_values(1.0) + [2.0, 0.2]
        ys = [0.0] * 7 + _edge_values(2.0) + [0.1] * 7 + [0.5, 0.1]
        on_edge = 5 + 5 + 5
    # an odd asphere keeps the lens off the reduced-algebra kernel; axis-parallel rays: the local x, y at the surface
    # are the launch values exactly, in both builds
    lens = _singlet(dict(radius=60.0, conic=0.0, surface_type="odd_asphere", coefficients=[0.0, 1.0e-4], aperture=ap))
    n = len(xs)
    rays_in = [np.array(xs), np.array(ys), np.full(n, -5.0), np.zeros(n), np.zeros(n), np.ones(n), np.ones(n), np.full(n, 0.55)]
    fast, strict = _both(lens, rays_in)
    a, b = fast.out.cpu().numpy(), strict.out.cpu().numpy()
    assert np.array_equal(a[6] == 0, b[6] == 0) and (b[6] == 0).any() and (b[6] != 0).any()
    assert np.array_equal(np.isnan(a), np.isnan(b))
    assert np.max(np.abs(a - b) / np.maximum(np.abs(b), 1.0)) <= 1e-10
    assert int(fast.fixups) >= on_edge          # the rays within 2^-36 of an edge went to the strict build
    assert int(fast.fixups) < n                 # ... the others did not


def test_adjacent_doubles_across_the_critical_angle():
    """Glass-to-air exit face tilted by 0.3 rad: the ray direction is swept, one double at a time, across the angle
    where the exit radicand 1 - u^2 (1 - dot^2) changes sign. 2001 adjacent values of M around the transition: the
    tolerance build's TIR pattern is the strict build's, ray for ray."""
    lens = _singlet(dict(radius=80.0), dict(radius=INF, rx=0.3))

    def dead(M):
        M = np.atleast_1d(np.asarray(M, dtype=np.float64))
        n = M.size
        rays_in = [np.zeros(n), np.zeros(n), np.full(n, -5.0), np.zeros(n), M, np.sqrt(1 - M * M), np.ones(n), np.full(n, 0.55)]
        fast, strict = _both(lens, rays_in)
        return np.isnan(fast.out[3].cpu().numpy()), np.isnan(strict.out[3].cpu().numpy()), fast

    lo, hi = 0.0, 0.9                         # refracted at lo, totally reflected at hi
    _, d, _ = dead([lo, hi])
    assert not d[0] and d[1]
    for _ in range(80):                        # bisection with the strict build down to adjacent doubles
        mid = 0.5 * (lo + hi)
        if mid == lo or mid == hi:
            break
        _, d, _ = dead([mid])
        lo, hi = (lo, mid) if d[0] else (mid, hi)
    bits = np.array([lo], dtype=np.float64).view(np.int64)[0]
    M = (bits + np.arange(-1000, 1001, dtype=np.int64)).view(np.float64)
    df, ds, fast = dead(M)
    assert ds.any() and not ds.all()
    assert np.array_equal(df, ds)
    assert int(fast.fixups) == M.size         # every one of them is inside the band


def test_paraxial_rays_on_a_paraboloid_take_the_strict_answer():
    """k = -1: a = L^2 + M^2 of the quadratic is ~1e-11 for rays a few microradians off the axis and the reference's
    (-b + sqrt d) / 2a cancels: its t depends on the evaluation order far beyond 1e-10 (nothing a tolerance could be
    stated for). Rays with |a| < 2^-20 are re-traced by the strict build, so the tolerance mode returns the reference's
    own values; rays exactly along the axis (a == 0 in both builds: the -c / 2b branch, standard.py:136-138) stay on
    the fast path."""
    lens = _singlet(dict(radius=40.0, conic=-1.0, dy=0.3), dict(radius=INF, rx=0.1))
    n = 4096
    rng = np.random.default_rng(1)
    x, y = rng.uniform(-3, 3, n), rng.uniform(-3, 3, n)
    L_ = rng.uniform(-4e-6, 4e-6, n)
    M_ = rng.uniform(-4e-6, 4e-6, n)
    L_[:64], M_[:64] = 0.0, 0.0
    rays_in = [x, y, np.full(n, -5.0), L_, M_, np.sqrt(1 - L_ * L_ - M_ * M_), np.ones(n), np.full(n, 0.55)]
    fast, strict = _both(lens, rays_in)
    a, b = fast.out.cpu().numpy(), strict.out.cpu().numpy()
    assert np.isfinite(b).all()
    # (a ray so close to the axis that L^2 + M^2 is absorbed by N^2 - N^2 has a == 0 exactly like the axial ones)
    assert n - 72 <= int(fast.fixups) <= n - 64
    same = (a == b).all(axis=0)
    assert int(same[64:].sum()) >= n - 72                              # re-traced: the strict build's bits
    assert np.max(np.abs(a - b) / np.maximum(np.abs(b), 1.0)) <= 1e-12


@pytest.mark.parametrize("tilted", [False, True], ids=["reduced_algebra_kernel", "interpreter_kernel"])
def test_tolerance_mode_near_the_axis_of_a_paraboloid(tilted):
    """Both tolerance-mode kernels on a paraboloid (k = -1), ray bundles within 1e-7 ... 0.1 rad of its axis: the
    leading coefficient a = L^2 + M^2 of the reference's quadratic goes to 0, its root cancels, and the reference's t
    carries rounding noise of ~4e-16 / a relative -- which only its own operation order reproduces. Rays with
    |a| < 2^-17 are re-traced by the strict build (band in trace_fx.cuh: fx_conic_t and trace_common.cuh:
    conic_distance), so tolerance mode stays within 1e-10 of the reference there too. Without the band the
    reduced-algebra kernel differs by 3.5 (relative) at 1e-7 rad and by 1e-9 at 1e-2 rad."""
    first = dict(radius=40.0, conic=-1.0, dy=0.3) if tilted else dict(radius=40.0, conic=-1.0)
    lens = _singlet(first, dict(radius=INF, rx=0.05) if tilted else None)
    packed = pack_surface_group(lens.surface_group.surfaces, [0.55])
    assert packed.fastkinds() != tilted
    n = 8192
    for amp in (1e-7, 1e-5, 1e-3, 1e-2, 0.1):
        rng = np.random.default_rng(1)
        x, y = rng.uniform(-3, 3, n), rng.uniform(-3, 3, n)
        L_, M_ = rng.uniform(-amp, amp, n), rng.uniform(-amp, amp, n)
        rays_in = [x, y, np.full(n, -5.0), L_, M_, np.sqrt(1 - L_ * L_ - M_ * M_), np.ones(n), np.full(n, 0.55)]
        dev = [torch.tensor(v, device="cuda") for v in rays_in]
        fast = fused_trace(packed, dev, fast=True)
        a, b = fast.out.cpu().numpy(), fused_trace(packed, dev, fast=False).out.cpu().numpy()
        assert np.isfinite(b).all()
        assert np.max(np.abs(a - b) / np.maximum(np.abs(b), 1.0)) <= 1e-10, amp
        if amp <= 1e-3:
            # a <= 2e-6: every ray -- but for those whose L^2 + M^2 is absorbed to an exact 0 (the -c / 2b branch)
            assert int(fast.fixups) >= n * 9 // 10, (amp, int(fast.fixups))
        elif amp >= 0.1:
            assert int(fast.fixups) < n // 100, (amp, int(fast.fixups))    # a handful that happen to be near the axis


@pytest.mark.parametrize("tilted", [False, True], ids=["reduced_algebra_kernel", "interpreter_kernel"])
def test_tolerance_mode_on_nearly_flat_surfaces(tilted):
    """A "flat" surface entered with a huge radius (R = 1e7 ... 1e10 instead of inf): b ~ N R in the reference's
    quadratic, (-b + sqrt d) cancels and its t carries ~5e-16 R of absolute rounding noise (1.7e-7 relative at
    R = 1e10, measured between the tolerance and the strict build before the band). Rays with |a| < 2^-20 |b| go to
    the strict build: within 1e-10 of the reference at every radius."""
    n = 4096
    rng = np.random.default_rng(3)
    x, y = rng.uniform(-3, 3, n), rng.uniform(-3, 3, n)
    L_, M_ = rng.uniform(-0.1, 0.1, n), rng.uniform(-0.1, 0.1, n)
    rays_in = [x, y, np.full(n, -5.0), L_, M_, np.sqrt(1 - L_ * L_ - M_ * M_), np.ones(n), np.full(n, 0.55)]
    dev = [torch.tensor(v, device="cuda") for v in rays_in]
    for R, k in ((1e3, 0.0), (1e5, -0.5), (1e7, 0.0), (1e8, -0.5), (1e10, 0.0)):
        lens = _singlet(dict(radius=R, conic=k, **(dict(rx=0.01) if tilted else {})), dict(radius=-R, conic=k))
        packed = pack_surface_group(lens.surface_group.surfaces, [0.55])
        assert packed.fastkinds() != tilted
        fast = fused_trace(packed, dev, fast=True)
        a, b = fast.out.cpu().numpy(), fused_trace(packed, dev, fast=False).out.cpu().numpy()
        assert np.isfinite(b).all()
        assert np.max(np.abs(a - b) / np.maximum(np.abs(b), 1.0)) <= 1e-10, R
        assert (int(fast.fixups) == n) == (R >= 1e7), (R, int(fast.fixups))
