"""BASELINE configs[1] at its FULL size (even-asphere double Gauss, 3 fields x 16 777 216 rays, fp64) through the
public API, checked with size-independent properties: the oracle cannot finish 50 M rays in seconds, so the
values themselves are pinned at small sizes (test_gpu_golden / fuzz / pupil) and here we check what must hold
for any batch: unit direction vectors, mirror symmetry of the lens, chunked == resident, fused spot sums == a
torch reduction of the outputs, tolerance mode == strict mode to 1e-10 with identical masks."""
import numpy as np
import pytest
import torch

from optiland_pr_b200 import backend as be
from optiland_pr_b200 import samples, spot
from optiland_pr_b200.trace import fused_trace

pytestmark = pytest.mark.gpu
NPF = 16_777_216
HX, HY = [0.0, 0.0, 0.0], [0.0, 10.0 / 14.0, 1.0]


@pytest.fixture(scope="module")
def traced():
    be.set_precision("float64")
    lens = samples.EvenAsphereDoubleGauss()
    lens.surface_group.record = "none"
    rng = np.random.default_rng(1)
    h = NPF // 2
    r, th = rng.uniform(size=h), rng.uniform(0, 2 * np.pi, size=h)
    px, py = np.sqrt(r) * np.cos(th), np.sqrt(r) * np.sin(th)
    # second half = mirror image of the first half about the y-z plane (all fields have Hx = 0)
    Px_h = torch.from_numpy(np.concatenate([px, -px])).pin_memory()
    Py_h = torch.from_numpy(np.concatenate([py, py])).pin_memory()
    rays = lens.trace(HX, HY, 0.5876, distribution=(Px_h, Py_h), spot=True)       # host grid: chunked, pipelined
    return lens, Px_h, Py_h, rays


def test_unit_directions_and_finite_outputs(traced):
    _, _, _, rays = traced
    assert len(rays) == 3 * NPF
    nrm = rays.L * rays.L + rays.M * rays.M + rays.N * rays.N
    assert float((nrm - 1.0).abs().max()) <= 1e-12
    assert bool(torch.isfinite(rays.x).all()) and bool(torch.isfinite(rays.opd).all())
    assert bool((rays.i > 0).all())


def test_mirror_symmetry_about_the_meridional_plane(traced):
    _, _, _, rays = traced
    h = NPF // 2
    for f in range(3):
        a = slice(f * NPF, f * NPF + h)
        b = slice(f * NPF + h, (f + 1) * NPF)
        assert float((rays.x[a] + rays.x[b]).abs().max()) <= 1e-12
        assert float((rays.y[a] - rays.y[b]).abs().max()) <= 1e-12
        assert float((rays.L[a] + rays.L[b]).abs().max()) <= 1e-13
        assert float((rays.opd[a] - rays.opd[b]).abs().max()) <= 1e-11


def test_chunked_host_grid_equals_resident_grid(traced):
    lens, Px_h, Py_h, rays = traced
    res = lens.trace(HX, HY, 0.5876, distribution=(Px_h.cuda(), Py_h.cuda()), spot=True)
    for name in ("x", "y", "z", "L", "M", "N", "i", "opd", "L0", "M0", "N0"):
        assert torch.equal(getattr(rays, name), getattr(res, name)), name
    # the raw sums are taken about each call's own shift (its chief rays); what they stand for must agree
    sa, sb = spot.spot_stats(rays.spot_sums), spot.spot_stats(res.spot_sums)
    for key in ("count", "cx", "cy", "rms", "intensity"):
        assert torch.allclose(sa[key], sb[key], rtol=1e-10, atol=1e-10), key


def test_fused_spot_sums_equal_a_reduction_of_the_outputs(traced):
    _, _, _, rays = traced
    rms = spot.spot_stats(rays.spot_sums)["rms"].cpu().numpy()
    x, y = rays.x.reshape(3, -1), rays.y.reshape(3, -1)
    ref = torch.sqrt(((x - x.mean(1, keepdim=True)) ** 2 + (y - y.mean(1, keepdim=True)) ** 2).mean(1)).cpu().numpy()
    assert np.allclose(rms, ref, rtol=1e-9, atol=0)
    assert np.array_equal(rays.spot_sums[:, 0].cpu().numpy(), np.full(3, float(NPF)))


def test_tolerance_mode_equals_strict_mode(traced):
    lens, Px_h, Py_h, rays = traced
    launched = lens.launch_rays(np.asarray(HX), np.asarray(HY), Px_h.cuda(), Py_h.cuda(), 0.5876)
    ins = launched.components()
    ins[7] = None
    # field by field: a multi-field Optic.trace gives every field its own Newton trip counts (optic.py), as separate
    # single-field calls of the reference would
    strict = torch.empty((8, 3 * NPF), dtype=torch.float64, device="cuda")
    for f in range(3):
        sl = slice(f * NPF, (f + 1) * NPF)
        packed = lens.surface_group.pack([0.5876])
        res = fused_trace(packed, [c[sl] if c is not None else None for c in ins], out=strict[:, sl], fast=False)
        assert [int(k) for k in res.nr_iters] == rays.nr_iters_per_field[f]
    out = torch.stack([rays.x, rays.y, rays.z, rays.L, rays.M, rays.N, rays.i, rays.opd])
    assert torch.equal(out.isnan(), strict.isnan())
    assert torch.equal(out[6] == 0, strict[6] == 0)
    err = ((out - strict).abs() / torch.clamp(strict.abs(), min=1.0)).max()
    assert float(err) <= 1e-10, float(err)
