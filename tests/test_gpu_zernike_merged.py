"""Tolerance-mode Zernike surfaces: the merged tables of the interpreter kernel (csrc/trace_geom.cuh: zernike_merge /
zernike_eval_merged -- per (|m|, cos / sin) ONE polynomial in rho^2, the reference's regularised slope quotients kept)
against the NumPy oracle's term-by-term evaluation (oracle/trace_np.py, zernike/base.py:42-299, geometries/zernike.py:
133-232): <= 1e-10 with identical masks and NaN patterns, on and next to the axis, with and without norm constants,
and the fall-back to the term loop for what the tables do not hold."""
import numpy as np
import pytest
import torch

from oracle import trace_np
from optiland_pr_b200 import backend as be
from optiland_pr_b200.optic import Optic
from optiland_pr_b200.prescription import pack_surface_group
from optiland_pr_b200.trace import fused_trace

pytestmark = pytest.mark.gpu
INF = float("inf")


def _lens(surfaces):
    be.set_precision("float64")
    lens = Optic("zernike")
    lens.add_surface(index=0, radius=INF, thickness=INF)
    for k, kw in enumerate(surfaces):
        lens.add_surface(index=1 + k, thickness=3.0, material="N-SF11" if k % 2 == 0 else "air", is_stop=(k == 0), **kw)
    lens.add_surface(index=1 + len(surfaces), radius=INF, thickness=20.0)
    lens.add_surface(index=2 + len(surfaces))
    lens.add_wavelength(0.55, is_primary=True)
    return lens


def _rays(n, seed, half=4.0):
    rng = np.random.default_rng(seed)
    x, y = rng.uniform(-half, half, n), rng.uniform(-half, half, n)
    # the axis and its neighbourhood: the reference's slopes carry 1 / (rho + 1e-14) and 1 / (rho^2 + 1e-14), which
    # change the answer by far more than 1e-10 for rho^2 <~ 1e-4 -- the merged form has to reproduce them, not the
    # exact gradient
    near = np.array([0.0, 1e-12, 1e-9, 1e-7, 1e-6, 1e-5, 1e-4, 1e-3, 1e-2, 0.1])
    x[: near.size], y[: near.size] = near, 0.0
    x[near.size: 2 * near.size], y[near.size: 2 * near.size] = -0.6 * near, 0.8 * near
    L = rng.uniform(-0.02, 0.02, n)
    M = rng.uniform(-0.02, 0.02, n)
    L[:2], M[:2] = 0.0, 0.0
    return [x, y, np.full(n, -5.0), L, M, np.sqrt(1 - L * L - M * M), np.ones(n), np.full(n, 0.55)]


def _compare(lens, rays_in, tol=1e-10):
    packed = pack_surface_group(lens.surface_group.surfaces, [0.55])
    with np.errstate(all="ignore"):
        want = np.array(trace_np.trace_packed(packed, trace_np.Rays(*rays_in))["records"])
    dev = [torch.tensor(rays_in[q], device="cuda") for q in range(8)]
    got = fused_trace(packed, dev, fast=True).record.cpu().numpy()
    strict = fused_trace(packed, dev, fast=False).record.cpu().numpy()
    assert np.array_equal(np.isnan(got), np.isnan(want))
    assert np.array_equal(got[:, 6] == 0, want[:, 6] == 0)
    m = np.isfinite(want)
    err = np.max(np.abs(got[m] - want[m]) / np.maximum(np.abs(want[m]), 1.0), initial=0.0)
    assert err <= tol, err
    return got, strict, want


@pytest.mark.parametrize("ztype", ["fringe", "noll", "standard"])
def test_thirty_six_terms_against_the_oracle(ztype):
    """36 coefficients of each index scheme (|m| up to 5 .. 7, radial polynomials up to rho^10; noll and standard carry
    norm constants, so the sag and the slopes use different tables), 4096 rays incl. the axis and its neighbourhood."""
    rng = np.random.default_rng(11)
    co = (rng.uniform(-1, 1, 36) * 2.0e-3).tolist()
    co[4] = 0.0                      # a zero coefficient: the reference skips it in the slopes only
    lens = _lens([dict(radius=45.0, conic=-0.4, surface_type="zernike", zernike_type=ztype, norm_radius=7.0,
                       coefficients=co, rx=0.01, dy=0.2)])
    got, strict, want = _compare(lens, _rays(4096, 3))
    assert np.isfinite(want).all()
    assert not np.array_equal(got, strict)      # the merged form is a different summation order: not the term loop


def test_slopes_next_to_the_axis_follow_the_regularised_quotients():
    """A pure tilt term Z = rho cos(phi): its exact x-slope is 1 / Rn everywhere, the reference's is
    (cos^2 rho / (rho + eps) + sin^2 rho^2 / (rho^2 + eps)) / Rn -- half of it at rho = 1e-7 on the y axis. The kernel's
    refracted direction there follows the reference, to 1e-10."""
    lens = _lens([dict(radius=INF, surface_type="zernike", zernike_type="fringe", norm_radius=2.0,
                       coefficients=[0.0, 5.0e-2])])
    n = 12
    rho = np.array([0.0, 1e-9, 3e-8, 1e-7, 3e-7, 1e-6, 1e-5, 1e-4, 1e-3, 1e-2, 0.1, 1.0]) * 2.0
    rays_in = [np.zeros(n), rho, np.full(n, -1.0), np.zeros(n), np.zeros(n), np.ones(n), np.ones(n), np.full(n, 0.55)]
    got, strict, want = _compare(lens, rays_in)
    # the regularisation is visible: the direction after the surface changes with rho although the surface is a plane
    assert abs(want[0, 3, 3] - want[0, 3, -1]) > 1e-4


def test_terms_the_tables_do_not_hold_keep_the_term_loop():
    """|m| = 9 (standard index 54: n = 9, m = 9) is past the tables (OLRT_ZT_M = 9): the surface stays on the term loop
    (a table indexed with it would be overrun and the answer wrong)."""
    co = [0.0] * 55
    co[3], co[12], co[54] = 1.0e-2, -3.0e-3, 2.0e-3
    lens = _lens([dict(radius=60.0, conic=0.0, surface_type="zernike", zernike_type="standard", norm_radius=7.0,
                       coefficients=co)])
    got, strict, want = _compare(lens, _rays(512, 5))
    assert np.isfinite(want).all()


def test_three_zernike_surfaces_two_tables():
    """OLRT_ZT_SURF = 2 tables per CTA: the third Zernike surface runs the term loop next to two merged ones."""
    rng = np.random.default_rng(2)
    sf = [dict(radius=r, conic=0.1, surface_type="zernike", zernike_type=t, norm_radius=8.0,
               coefficients=(rng.uniform(-1, 1, 15) * 1.0e-3).tolist())
          for r, t in ((50.0, "fringe"), (-70.0, "noll"), (90.0, "standard"))]
    _compare(_lens(sf), _rays(1024, 7))


def test_prescription_batch_rebuilds_the_tables():
    """Two prescriptions in one launch (coefficients differ): every CTA rebuilds its tables when its tile moves to the
    next prescription -- same rays as two single launches, bit for bit."""
    co_a = [0.0, 1.0e-2, -2.0e-2, 1.0e-2, 0.0, 3.0e-3, 4.0e-3, -2.0e-3]
    co_b = [1.0e-3, -1.0e-2, 1.0e-2, 2.0e-2, 1.0e-3, 0.0, -4.0e-3, 2.0e-3]
    kw = dict(radius=50.0, conic=0.1, surface_type="zernike", zernike_type="noll", norm_radius=8.0)
    la, lb = _lens([dict(coefficients=co_a, **kw)]), _lens([dict(coefficients=co_b, **kw)])
    pa = pack_surface_group(la.surface_group.surfaces, [0.55])
    pb = pack_surface_group(lb.surface_group.surfaces, [0.55])
    n = 1024
    rays = [torch.tensor(v, device="cuda") for v in _rays(n, 9)]
    one_a = fused_trace(pa, rays, fast=True).out
    one_b = fused_trace(pb, rays, fast=True).out
    presc = torch.tensor(np.concatenate([pa.presc, pb.presc], axis=0), device="cuda")
    both = fused_trace(pa, [torch.cat([v, v]) for v in rays], presc=presc, n_presc=2, rays_per_presc=n, fast=True).out
    assert torch.equal(both[:, :n], one_a) and torch.equal(both[:, n:], one_b)
    assert not torch.equal(one_a, one_b)


@pytest.mark.parametrize("ztype", ["fringe", "noll", "standard"])
def test_plate_masks_equal_the_strict_build_on_two_million_rays(ztype):
    """Bulk check next to tests/test_gpu_interpreter_margins.py (the interpreter's tolerance build
    takes its clip / TIR / miss decisions with arithmetic that differs from the strict build's by rounding, and hands
    a ray within 2^-36 of a threshold to the strict build). On the C3 plate (annular, elliptical, rectangular and
    boolean apertures, TIR on part of the pupil), both fields, 2 M random rays each: not one ray decides differently,
    and the values agree to 1e-10."""
    from optiland_pr_b200 import samples
    from optiland_pr_b200.optic import random_pupil_device
    be.set_device("cuda:0")
    be.set_precision("float64")
    lens = samples.FreeformPlate(ztype)
    lens.surface_group.record = "none"
    wl = lens.primary_wavelength
    n = 1 << 21
    dev = torch.device("cuda:0")
    Px, Py = random_pupil_device(n, 2, dev)
    packed = lens.surface_group.pack([float(wl)])
    for f0 in lens.get_field_coords():
        rays = lens.launch_rays(np.array([f0[0]]), np.array([f0[1]]), Px, Py, wl)
        ins = rays.components()
        ins[7] = None
        a = fused_trace(packed, ins, fast=True).out
        b = fused_trace(packed, ins, fast=False).out
        assert bool((torch.isnan(a) == torch.isnan(b)).all())
        assert bool(((a[6] == 0) == (b[6] == 0)).all())
        assert int((b[6] == 0).sum()) > 0                                   # the masks are not trivial
        d = (torch.nan_to_num(a) - torch.nan_to_num(b)).abs() / torch.nan_to_num(b).abs().clamp(min=1.0)
        assert float(d.max()) <= 1e-10
