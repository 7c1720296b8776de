"""Reduced-algebra adjoint kernel (csrc/trace_bx.cuh, the default backward for plane / conic / even-asphere
lenses) against the general adjoint kernel (csrc/trace_bwd.cuh, pinned to the reference's torch autograd by
tests/test_gpu_grads.py) on the same rays: both differentiate Surface.trace
(optiland/surfaces/standard_surface.py:186-233); the two formulations share no intermediate algebra, so agreement
to 1e-9 on thousands of rays -- vignetted, absorbed, reflected, decentred, launched with zero intensity, three
wavelengths, gradients entering at every recorded surface -- checks both."""
import numpy as np
import pytest
import torch

from optiland_pr_b200 import samples
from optiland_pr_b200.apertures import RadialAperture
from optiland_pr_b200.autograd import trace_differentiable
from optiland_pr_b200.optic import Optic
from optiland_pr_b200.prescription import pack_surface_group

pytestmark = pytest.mark.gpu
INF = float("inf")


@pytest.fixture(autouse=True)
def _fp64():
    from optiland_pr_b200 import backend as be
    be.set_device("cuda:0")
    be.set_precision("float64")
    yield


class _Mixed(Optic):
    """decentred conic with a radial aperture, absorbing glass, K = 2..3 even asphere with a conic constant,
    a conic mirror, a decentred plane with an annular aperture, plane image"""

    def __init__(self):
        super().__init__("mixed")
        self.add_surface(index=0, radius=INF, thickness=INF)
        self.add_surface(index=1, radius=40.0, conic=-0.6, thickness=6.0, material="N-SF11", dx=0.3, dy=-0.2,
                         aperture=RadialAperture(r_max=9.0))
        self.add_surface(index=2, radius=-55.0, conic=0.4, thickness=4.0, surface_type="even_asphere",
                         coefficients=[1.0e-4, -3.0e-6, 2.0e-8, -1.0e-10, 2.0e-13], is_stop=True,
                         aperture=RadialAperture(r_max=9.5))
        self.add_surface(index=3, radius=INF, thickness=30.0, material=("F2", "schott"), dx=-0.1,
                         aperture=RadialAperture(r_max=9.0, r_min=0.5))
        self.add_surface(index=4, radius=-120.0, conic=-1.3, thickness=-25.0, material="mirror")
        self.add_surface(index=5, radius=80.0, thickness=-3.0, material="SK16", surface_type="even_asphere",
                         conic=0.0, coefficients=[0.0, 2.0e-6])
        self.add_surface(index=6, radius=INF, thickness=-10.0)
        self.add_surface(index=7)
        self.set_aperture("EPD", 16.0)
        self.set_field_type("angle")
        self.add_field(y=0)
        self.add_field(y=3)
        self.add_wavelength(0.48)
        self.add_wavelength(0.55, is_primary=True)
        self.add_wavelength(0.65)
        self.set_paraxial(EPL=0.0, EPD=16.0)


def _both(lens, wavelengths, n, seed, use_record_grads=True, zero_intensity=True):
    rng = np.random.default_rng(seed)
    r_, th = np.sqrt(rng.uniform(size=n)), rng.uniform(0, 2 * np.pi, n)
    wl = np.asarray(wavelengths)[rng.integers(0, len(wavelengths), n)]
    Hy = rng.choice([0.0, 0.7, 1.0], n)
    rays = lens.generate_rays(torch.zeros(n, dtype=torch.float64, device="cuda"),
                              torch.from_numpy(Hy).cuda(), torch.from_numpy(r_ * np.cos(th)).cuda(),
                              torch.from_numpy(r_ * np.sin(th)).cuda(), torch.from_numpy(wl).cuda())
    ins = rays.components()
    if len(wavelengths) == 1:
        ins[7] = None
    if zero_intensity:
        ins[6] = ins[6].clone()
        ins[6][::17] = 0.0
        ins[6][1::17] = 0.5
    pk = pack_surface_group(lens.surface_group.surfaces, list(wavelengths), record="all")
    S = pk.n_surfaces
    gout = torch.from_numpy(rng.normal(size=(8, n))).cuda()
    grec = torch.from_numpy(rng.normal(size=(S, 8, n))).cuda() if use_record_grads else None
    res = []
    for generic in (True, False):
        presc = torch.tensor(pk.presc[0], dtype=torch.float64, device="cuda", requires_grad=True)
        leaf = [t.detach().clone().requires_grad_(q < 7) if t is not None else None for q, t in enumerate(ins)]
        opd_in = torch.zeros(n, dtype=torch.float64, device="cuda", requires_grad=True)
        out, record, _ = trace_differentiable(pk, presc, leaf, opd_in=opd_in, fast=True, generic_adjoint=generic)
        loss = (out * gout).nan_to_num(0.0).sum()
        if grec is not None:
            loss = loss + (record * grec).nan_to_num(0.0).sum()
        loss.backward()
        gin = torch.stack([leaf[q].grad for q in range(7)] + [opd_in.grad]).cpu().numpy()
        res.append((gin, presc.grad.cpu().numpy(), out.detach().cpu().numpy(), pk.nr_iters()))
    return res


def _check(res, tol=1e-9):
    (g0, p0, out, iters), (g1, p1, _, _) = res
    ok = np.isfinite(out).all(axis=0)          # rays the forward completed (misses / TIR are NaN in both)
    assert ok.mean() > 0.5
    assert np.array_equal(np.isnan(g0), np.isnan(g1))
    for q in range(8):
        a, b = g1[q][ok], g0[q][ok]
        scale = np.abs(b).max() + 1e-300
        err = np.abs(a - b).max() / scale
        assert err <= tol, f"ray gradient {q}: {err}"
    if np.isfinite(p0).all():
        scale = np.abs(p0).max()
        nz = p0 != 0
        assert np.array_equal(nz, p1 != 0)
        assert np.abs(p1 - p0).max() <= 1e-8 * scale, np.stack([p1[nz], p0[nz]], 1)
    return iters


@pytest.mark.parametrize("name", ["CookeTriplet", "DoubleGauss", "EvenAsphereDoubleGauss", "AsphericSinglet"])
def test_reduced_adjoint_matches_general_adjoint_on_sample_lenses(name):
    lens = getattr(samples, name)()
    wls = lens.wavelengths if len(lens.wavelengths) > 1 else [lens.wavelengths[0]]
    iters = _check(_both(lens, wls, 20_000, seed=3))
    if name == "EvenAsphereDoubleGauss":
        assert max(iters) >= 2      # surface 5 takes two Newton updates: the unrolled adjoint with a stored iterate


def test_reduced_adjoint_mixed_lens_apertures_mirror_decenters():
    lens = _Mixed()
    res = _both(lens, lens.wavelengths, 30_000, seed=5)
    iters = _check(res)
    assert max(iters) >= 1
    out = res[0][2]
    assert (out[6] == 0).any() and (out[6] > 0).any()      # some rays vignetted


def test_reduced_adjoint_valid_rays_only_prescription_gradient():
    """no NaN rays (paraxial bundle): prescription gradients are finite and agree entry by entry"""
    lens = samples.EvenAsphereDoubleGauss()
    n = 50_000
    rng = np.random.default_rng(11)
    r_, th = 0.8 * np.sqrt(rng.uniform(size=n)), rng.uniform(0, 2 * np.pi, n)
    Px, Py = torch.from_numpy(r_ * np.cos(th)).cuda(), torch.from_numpy(r_ * np.sin(th)).cuda()
    rays = lens.generate_rays(torch.zeros_like(Px), torch.full_like(Px, 0.7), Px, Py, 0.5876)
    ins = rays.components()
    ins[7] = None
    pk = pack_surface_group(lens.surface_group.surfaces, [0.5876], record="all")
    got = []
    for generic in (True, False):
        presc = torch.tensor(pk.presc[0], dtype=torch.float64, device="cuda", requires_grad=True)
        out, record, _ = trace_differentiable(pk, presc, ins, fast=True, generic_adjoint=generic)
        x, y = out[0], out[1]
        loss = ((x - x.mean()) ** 2 + (y - y.mean()) ** 2).mean() + 1e-3 * out[7].mean() + out[6].mean()
        loss.backward()
        got.append(presc.grad.cpu().numpy())
    p0, p1 = got
    assert np.isfinite(p0).all() and np.isfinite(p1).all() and (p0 != 0).sum() > 30
    nz = p0 != 0
    rel = np.abs(p1[nz] - p0[nz]) / np.maximum(np.abs(p0[nz]), 1e-6 * np.abs(p0).max())
    assert rel.max() <= 1e-8, rel.max()


@pytest.mark.parametrize("name,n", [("EvenAsphereDoubleGauss", 70_001), ("CookeTriplet", 4_096), ("_Mixed", 33_000)])
def test_blocked_checkpoints_give_the_same_gradients(name, n):
    """Tile-blocked checkpoints (OLRT_FWD_CKPT_BLOCKED / OLRT_BWD_CKPT_BLOCKED: 6 components per surface, one bulk copy
    per tile and surface in the backward) against the [S, 8, N] records: same forward state, same adjoint arithmetic,
    so the gradients agree to rounding of the accumulation order; ragged last tile included."""
    lens = _Mixed() if name == "_Mixed" else getattr(samples, name)()
    wls = list(lens.wavelengths)
    rng = np.random.default_rng(17)
    r_, th = np.sqrt(rng.uniform(size=n)), rng.uniform(0, 2 * np.pi, n)
    wl = np.asarray(wls)[rng.integers(0, len(wls), n)]
    Hy = rng.choice([0.0, 0.7, 1.0], n)
    rays = lens.generate_rays(torch.zeros(n, dtype=torch.float64, device="cuda"), torch.from_numpy(Hy).cuda(),
                              torch.from_numpy(r_ * np.cos(th)).cuda(), torch.from_numpy(r_ * np.sin(th)).cuda(),
                              torch.from_numpy(wl).cuda())
    ins = rays.components()
    ins[6] = ins[6].clone()
    ins[6][::19] = 0.0
    gout = torch.from_numpy(rng.normal(size=(8, n))).cuda()
    got = []
    for blocked in (False, True):
        pk = pack_surface_group(lens.surface_group.surfaces, wls, record="none" if blocked else "all")
        presc = torch.tensor(pk.presc[0], dtype=torch.float64, device="cuda", requires_grad=True)
        leaf = [t.detach().clone().requires_grad_(q < 7) for q, t in enumerate(ins)]
        opd_in = torch.zeros(n, dtype=torch.float64, device="cuda", requires_grad=True)
        out, record, dir0 = trace_differentiable(pk, presc, leaf, opd_in=opd_in, fast=True, blocked=blocked)
        assert (record.numel() == 0) == blocked
        (out * gout).nan_to_num(0.0).sum().backward()
        gin = torch.stack([leaf[q].grad for q in range(7)] + [opd_in.grad]).cpu().numpy()
        got.append((out.detach().cpu().numpy(), dir0.cpu().numpy(), gin, presc.grad.cpu().numpy()))
    (o0, d0, g0, p0), (o1, d1, g1, p1) = got
    assert np.array_equal(o0, o1, equal_nan=True) and np.array_equal(d0, d1, equal_nan=True)
    # per-ray gradients: the same arithmetic on the same states, except that the blocked layout keeps no N -- the
    # backward rebuilds it as sqrt(1 - L^2 - M^2) (include/olrt.h), a few ulps from the stored one
    assert np.array_equal(np.isnan(g0), np.isnan(g1))
    for q in range(8):
        fin = np.isfinite(g0[q])
        assert np.abs(g1[q][fin] - g0[q][fin]).max() <= 1e-12 * max(np.abs(g0[q][fin]).max(), 1e-300), q
    if np.isfinite(p0).all():
        assert np.array_equal(p0 != 0, p1 != 0)
        assert np.abs(p1 - p0).max() <= 1e-11 * np.abs(p0).max()
    else:
        assert np.array_equal(np.isnan(p0), np.isnan(p1))


def test_public_api_gradients_do_not_depend_on_the_checkpoint_layout():
    """Optic.trace with record 'none' (tile-blocked checkpoints) and record 'all' ([S, 8, N] records): the same
    d(merit) / d(radius, thickness-driven frame z) through the public autograd path."""
    from optiland_pr_b200.autograd import _rms_merit
    rng = np.random.default_rng(4)
    rr, th = np.sqrt(rng.uniform(size=3000)), rng.uniform(0, 2 * np.pi, 3000)
    Px, Py = rr * np.cos(th), rr * np.sin(th)
    fields = [(0.0, 0.0), (0.0, 0.7), (0.0, 1.0)]
    grads = []
    for mode in ("all", "none"):
        lens = samples.CookeTriplet()
        lens.surface_group.record = mode
        r1 = torch.tensor(22.01359, dtype=torch.float64, device="cuda", requires_grad=True)
        r4 = torch.tensor(20.29192, dtype=torch.float64, device="cuda", requires_grad=True)
        lens.set_radius(r1, 1)
        lens.set_radius(r4, 4)
        loss = _rms_merit(lens, fields, 0.55, Px, Py)
        loss.backward()
        grads.append((float(loss), float(r1.grad), float(r4.grad)))
        used_blocked = lens.surface_group.last_trace.record is None
        assert used_blocked == (mode == "none")
    (l0, a0, b0), (l1, a1, b1) = grads
    assert l0 == l1
    assert abs(a1 - a0) <= 1e-11 * abs(a0) and abs(b1 - b0) <= 1e-11 * abs(b0)


@pytest.mark.parametrize("blocked", [False, True])
def test_fp32_gradients_track_fp64(blocked):
    """olrt_trace_bwd_f32 (same kernels instantiated for float, both checkpoint layouts): the merit gradient w.r.t. the
    prescription agrees with the fp64 one to fp32 accuracy on a well-conditioned (paraxial) bundle."""
    lens = samples.EvenAsphereDoubleGauss()
    n = 20_000
    rng = np.random.default_rng(23)
    r_, th = 0.7 * np.sqrt(rng.uniform(size=n)), rng.uniform(0, 2 * np.pi, n)
    Px, Py = torch.from_numpy(r_ * np.cos(th)).cuda(), torch.from_numpy(r_ * np.sin(th)).cuda()
    rays = lens.generate_rays(torch.zeros_like(Px), torch.full_like(Px, 0.7), Px, Py, 0.5876)
    ins64 = rays.components()
    ins64[7] = None
    pk = pack_surface_group(lens.surface_group.surfaces, [0.5876], record="none" if blocked else "all")
    got = {}
    for dt in (torch.float64, torch.float32):
        ins = [t.to(dt) if t is not None else None for t in ins64]
        presc = torch.tensor(pk.presc[0], dtype=dt, device="cuda", requires_grad=True)
        try:
            out, _, _ = trace_differentiable(pk, presc, ins, fast=True, blocked=blocked)
        except NotImplementedError as e:
            # fp32 cannot push the asphere residual below tol = 1e-6 in 3 updates: the blocked layout (reduced-algebra
            # adjoint only) refuses, SurfaceGroup.trace then falls back to records + the general adjoint kernel
            assert blocked and dt == torch.float32 and "Newton" in str(e)
            return
        x, y = out[0].double(), out[1].double()
        loss = ((x - x.mean()) ** 2 + (y - y.mean()) ** 2).mean()
        loss.backward()
        got[dt] = (float(loss), presc.grad.double().cpu().numpy())
    l64, g64 = got[torch.float64]
    l32, g32 = got[torch.float32]
    assert abs(l32 - l64) <= 2e-3 * abs(l64)
    nz = g64 != 0
    assert nz.sum() >= 30 and np.array_equal(nz, g32 != 0)
    # asphere-coefficient gradients are orders of magnitude above the rest: every entry relative to itself, with a
    # floor for the sums of cancelling per-ray terms
    err = np.abs(g32[nz] - g64[nz]) / (np.abs(g64[nz]) + 1e-4 * np.median(np.abs(g64[nz])))
    assert err.max() <= 5e-2, err.max()
