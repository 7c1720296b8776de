"""The drop-in on the UNMODIFIED reference, on the GPU.

Every test runs a workload of tests/ref_cases.py twice through the reference's own classes (imported from the verbatim
copy under oracle/_ref on the GPU box): once on the stock path -- NumPy backend (fp64 oracle of record), or torch CPU
fp64 for gradients -- and once with `be.set_backend("torch"); be.set_device("cuda")` and `dropin.install()`, where
`SurfaceGroup.trace` is ONE launch of the fused kernels. `dropin.stats()` proves which path ran.
Bars (BASELINE.json north_star): fp64 <= 1e-10 relative, identical vignetting masks; fp32 <= 1e-5; gradients <= 1e-8.
The reference's own goldens (tests/test_analysis.py:69-100, tests/test_operand.py:265-325, tests/test_wavefront.py:
135-140) are asserted at the reference's tolerance (rtol 1e-5, tests/utils.py:6-13)."""
import numpy as np
import pytest
import torch

import ref_cases as rc
from optiland_pr_b200 import dropin

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _clean():
    yield
    dropin.uninstall()
    rc.restore()


def _rel(a, b, floor=1.0):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, (a.shape, b.shape)
    assert np.array_equal(np.isnan(a), np.isnan(b)), "NaN pattern differs"
    m = ~np.isnan(b)
    if not m.any():
        return 0.0
    return float(np.max(np.abs(a[m] - b[m]) / np.maximum(np.abs(b[m]), floor)))


def _fused(precision="float64", grad=False, **kw):
    rc.use_torch("cuda", precision, grad=grad)
    dropin.install(**kw)


def _only_fused(st, grad=False):
    """Every RealRays trace went through the kernels (paraxial traces are the one allowed fall-through)."""
    bad = {k: v for k, v in st.items() if k.startswith("stock:") and k != "stock:not RealRays"}
    assert not bad, st
    assert st.get("fused_grad" if grad else "fused", 0) > 0, st


# reference goldens, tests/test_analysis.py:69-100 (CookeTriplet of optiland.samples.objectives)
GEO = [[0.00597244087781, 0.00628645771124, 0.00931911440064],
       [0.03717783072826, 0.03864613392848, 0.04561512437816],
       [0.01951655430245, 0.02342659090311, 0.03747033587405]]
RMS = [[0.003791335461448, 0.004293689564257, 0.006195618755672],
       [0.015694600107671, 0.016786721284464, 0.019109151416248],
       [0.013229165357157, 0.012081348897953, 0.013596802321537]]


@pytest.mark.parametrize("fast_math", [True, False], ids=["tolerance", "strict"])
def test_spot_diagram_of_the_reference_on_the_fused_trace(fast_math):
    rc.use_numpy()
    want = rc.spot_diagram("cooke")
    _fused(fast_math=fast_math)
    got = rc.spot_diagram("cooke")
    _only_fused(dropin.stats())
    np.testing.assert_allclose(got["rms"], RMS, rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(got["geo"], GEO, rtol=1e-5, atol=1e-7)
    for k in ("rms", "geo", "centroid"):
        # radii of ~1e-2 mm from image coordinates of ~20 mm known to 1e-10 relative
        assert np.max(np.abs(got[k] - want[k])) <= 1e-10 * 20.0, k


@pytest.mark.parametrize("name", ["double_gauss", "cooke_stopped", "hubble", "reverse_telephoto"])
def test_spot_diagram_other_lenses(name):
    rc.use_numpy()
    want = rc.spot_diagram(name)
    _fused()
    got = rc.spot_diagram(name)
    _only_fused(dropin.stats())
    for k in ("rms", "geo", "centroid"):
        assert np.max(np.abs(got[k] - want[k])) <= 1e-10 * max(1.0, float(np.max(np.abs(want["centroid"])))), k


def test_rms_spot_size_operand():
    """RayOperand.rms_spot_size reads surface_group.x[-1, :] (operand/ray.py:336-340); reference golden
    tests/test_operand.py:277-288."""
    rc.use_numpy()
    want = [rc.rms_spot_operand("hubble", 0.0, 1.0, 0.55, 100), rc.rms_spot_operand("cooke", 0.0, 0.7, "all", 12),
            rc.rms_spot_operand("cooke", 0.3, 1.0, 0.48, 5000, "random")]
    _fused()
    got = [rc.rms_spot_operand("hubble", 0.0, 1.0, 0.55, 100), rc.rms_spot_operand("cooke", 0.0, 0.7, "all", 12)]
    _only_fused(dropin.stats())
    np.testing.assert_allclose(got[0], 0.025626727777956947, rtol=1e-5)
    for g, w in zip(got, want):
        assert abs(g - w) <= 1e-10 * 20.0
    # "random": the torch backend draws its own pupil (torch RNG != NumPy RNG): statistically the same spot
    g = rc.rms_spot_operand("cooke", 0.3, 1.0, 0.48, 5000, "random")
    assert abs(g - want[2]) <= 0.05 * want[2]


def test_single_ray_operands_read_positive_surface_indices():
    rc.use_numpy()
    want = rc.ray_operands("cooke")
    _fused()
    got = rc.ray_operands("cooke")
    _only_fused(dropin.stats())
    for k in want:
        assert abs(got[k] - want[k]) <= 1e-10 * max(1.0, abs(want[k])), k


def test_opd_difference_operand_and_wavefront():
    """Wavefront.get_data / ChiefRayStrategy (wavefront/strategy.py:176-229) and RayOperand.OPD_difference
    (operand/ray.py:342-389; goldens tests/test_operand.py:293-325). OPD in waves is a difference of long paths:
    1e-10 relative on the paths is 1e-10 * path / wavelength in waves -- the bar for tolerance mode. On the Hubble the
    reference's own quadratic (standard.py:105-134) is ill-conditioned at the nearly parabolic primary (a = 1 + k N^2 =
    -0.0023: (-b - sqrt(d)) / 2a cancels ~500x), so its per-ray OPD carries ~1e-5 waves of rounding noise; only the
    strict kernel, which rounds where the reference rounds, is held to the reference's goldens at its rtol 1e-5."""
    rc.use_numpy()
    want_d = [rc.opd_difference("hubble", 0.0, 1.0, 3), rc.opd_difference("hubble", 0.0, 0.0, 3),
              rc.opd_difference("hubble", 0.0, 1.0, 100, "hexapolar")]
    want_w = rc.wavefront("cooke")
    want_w2 = rc.wavefront("double_gauss", field=(0.0, 0.7), wl=0.5876)
    hubble_path_waves = 2.0e4 / 0.55e-3
    for fast_math in (False, True):
        _fused(fast_math=fast_math)
        got_d = [rc.opd_difference("hubble", 0.0, 1.0, 3), rc.opd_difference("hubble", 0.0, 0.0, 3),
                 rc.opd_difference("hubble", 0.0, 1.0, 100, "hexapolar")]
        got_w = rc.wavefront("cooke")
        got_w2 = rc.wavefront("double_gauss", field=(0.0, 0.7), wl=0.5876)
        _only_fused(dropin.stats())
        if not fast_math:
            np.testing.assert_allclose(got_d, [0.0732334055041471, 0.0013295118803701674, 0.2211995620762635],
                                       rtol=1e-5)
        for g, w in zip(got_d, want_d):
            assert abs(g - w) <= 1e-10 * hubble_path_waves
        for got, want, path in ((got_w, want_w, 70.0 / 0.55e-3), (got_w2, want_w2, 200.0 / 0.5876e-3)):
            assert np.max(np.abs(got["opd"] - want["opd"])) <= 1e-10 * path
            for k in ("pupil_x", "pupil_y", "pupil_z", "intensity"):
                assert _rel(got[k], want[k]) <= 1e-10, k
            assert abs(got["radius"] - want["radius"]) <= 1e-10 * abs(want["radius"])
        dropin.uninstall()


def _check_records(got, want):
    """fp64 bar: positions <= 1e-10 of the bundle's largest coordinate at that surface (relative error of the position
    vectors, never looser than 1e-10 mm), direction cosines <= 1e-10, OPD <= 1e-10 relative; masks bit-identical."""
    assert np.array_equal(got["intensity"] == 0, want["intensity"] == 0)
    scale = np.maximum(1.0, np.nanmax(np.abs(np.stack([want["x"], want["y"], want["z"]])), axis=(0, 2)))
    for k in ("x", "y", "z"):
        assert np.array_equal(np.isnan(got[k]), np.isnan(want[k])), k
        assert np.nanmax(np.abs(got[k] - want[k]) / scale[:, None], initial=0.0) <= 1e-10, k
    for k in ("rays.x", "rays.y", "rays.z"):
        assert np.nanmax(np.abs(got[k] - want[k]), initial=0.0) <= 1e-10 * scale[-1], k
    for k in ("L", "M", "N", "intensity", "opd", "rays.L", "rays.M", "rays.N", "rays.i", "rays.opd", "rays.L0",
              "rays.M0", "rays.N0"):
        assert _rel(got[k], want[k]) <= 1e-10, k


@pytest.mark.parametrize("name", ["cooke", "cooke_stopped", "double_gauss", "aspheric_singlet", "hubble",
                                  "reverse_telephoto"])
def test_every_recorded_surface_matches_the_numpy_reference(name):
    """surface_group.x[k, :] ... for every k (record="all"), the returned rays and L0 M0 N0 against the reference on
    its NumPy backend; tolerance mode (the default) and strict mode."""
    args = (0.0, 1.0, 0.55, 24, "hexapolar")
    rc.use_numpy()
    want = rc.trace_records(name, *args)
    for fast_math in (True, False):
        _fused(fast_math=fast_math)
        got = rc.trace_records(name, *args)
        _only_fused(dropin.stats())
        if name == "cooke_stopped":
            assert (want["intensity"][-1] == 0).any() and (want["intensity"][-1] != 0).any()
        _check_records(got, want)


@pytest.mark.parametrize("name", ["cooke", "cooke_stopped", "double_gauss", "hubble", "reverse_telephoto"])
def test_strict_mode_gives_the_bits_of_the_reference_torch_path(name):
    """Same inputs, bit for bit: the launch rays are made by the reference's own torch ops on the GPU in both runs;
    the stock loop (torch-eager CUDA kernels, IEEE + - * / sqrt, one rounding per op) and the strict fused kernel
    must then produce identical x y z L M N opd on closed-form lenses (intensity goes through exp: <= 1e-12)."""
    args = (0.3, 0.7, 0.65, 24, "hexapolar")
    rc.use_torch("cuda", "float64")
    want = rc.trace_records(name, *args)
    _fused(fast_math=False)
    got = rc.trace_records(name, *args)
    _only_fused(dropin.stats())
    for k in ("x", "y", "z", "L", "M", "N", "opd", "rays.x", "rays.y", "rays.L0", "rays.M0", "rays.N0"):
        assert np.array_equal(got[k], want[k], equal_nan=True), k
    assert np.array_equal(got["intensity"] == 0, want["intensity"] == 0)
    assert _rel(got["intensity"], want["intensity"]) <= 1e-12


def test_trace_generic_single_chief_ray_and_arrays():
    rc.use_numpy()
    want1 = rc.trace_generic("cooke", 0.0, 1.0, 0.0, 0.0, 0.55)
    Px, Py = np.linspace(-0.9, 0.9, 7), np.linspace(0.8, -0.8, 7)
    want7 = rc.trace_generic("cooke", np.zeros(7), np.full(7, 0.7), Px, Py, 0.65)
    _fused()
    got1 = rc.trace_generic("cooke", 0.0, 1.0, 0.0, 0.0, 0.55)
    got7 = rc.trace_generic("cooke", np.zeros(7), np.full(7, 0.7), Px, Py, 0.65)
    _only_fused(dropin.stats())
    assert got1["x"].shape == want1["x"].shape == (8, 1)
    for got, want in ((got1, want1), (got7, want7)):
        for k in want:
            assert _rel(got[k], want[k]) <= 1e-10, k


def test_fp32_precision_runs_the_fp32_kernel():
    args = (0.0, 0.7, 0.55, 24, "hexapolar")
    rc.use_numpy()
    want = rc.trace_records("cooke", *args)
    _fused(precision="float32")
    got = rc.trace_records("cooke", *args)
    _only_fused(dropin.stats())
    assert np.array_equal(got["intensity"] == 0, want["intensity"] == 0)
    for k in ("x", "y", "z"):
        assert np.max(np.abs(got[k] - want[k])) <= 1e-5 * max(1.0, np.max(np.abs(want[k]))), k
    for k in ("L", "M", "N"):
        assert np.max(np.abs(got[k] - want[k])) <= 1e-5, k
    assert _rel(got["opd"], want["opd"]) <= 1e-5


def test_record_last_keeps_the_image_surface_row():
    rc.use_numpy()
    want = rc.trace_records("cooke", 0.0, 1.0, 0.55, 12, "hexapolar")
    _fused(record="last")
    got = rc.trace_records("cooke", 0.0, 1.0, 0.55, 12, "hexapolar")
    _only_fused(dropin.stats())
    assert got["x"].shape[0] == 2          # the object surface's launch state + the image surface
    for k in ("x", "y", "z", "L", "M", "N", "intensity", "opd"):
        assert _rel(got[k][-1], want[k][-1]) <= 1e-10, k


class _IdentityBSDF:
    """A BSDF (interactions/base.py:111-118 calls `scatter`) that leaves the rays alone: outside the kernels."""

    def scatter(self, rays, nx, ny, nz):
        return rays


def test_surfaces_outside_the_kernels_fall_through_to_the_stock_loop():
    """NotEligible (a BSDF here) -> the reference's own loop on CUDA tensors: the bits of the unpatched reference."""
    def run():
        lens = rc.make_lens("cooke")
        lens.surface_group.surfaces[2].interaction_model.bsdf = _IdentityBSDF()
        rays = lens.trace(0.0, 1.0, 0.55, 6, "hexapolar")
        return rc.to_np(rays.x), rc.to_np(rays.i), rc.to_np(lens.surface_group.y)
    rc.use_torch("cuda", "float64")
    want = run()
    dropin.install()
    got = run()
    st = dropin.stats()
    assert any(k.startswith("stock:NotEligible") for k in st) and "fused" not in st, st
    for g, w in zip(got, want):
        assert np.array_equal(g, w, equal_nan=True)


def test_polarized_rays_fall_through():
    """PolarizedRays is a RealRays subclass the kernels do not implement (Fresnel coatings need it)."""
    from optiland.rays import PolarizationState
    rc.use_torch("cuda", "float64")
    dropin.install()
    lens = rc.make_lens("cooke")
    lens.set_polarization(PolarizationState(is_polarized=False))
    lens.surface_group.set_fresnel_coatings()
    rays = lens.trace(0.0, 1.0, 0.55, 6, "hexapolar")
    st = dropin.stats()
    assert st.get("stock:not RealRays", 0) >= 1 and "fused" not in st, st
    assert np.isfinite(rc.to_np(rays.i)).all() and (rc.to_np(rays.i) < 1.0).all()


@pytest.mark.parametrize("name", ["cooke", "aspheric_singlet"])
def test_torch_adam_on_the_fused_trace_matches_the_reference_autograd(name):
    """One TorchAdamOptimizer run (optimization/optimizer/torch/base.py:95-154) on RMS-spot operands with radius,
    thickness, conic and asphere-coefficient variables: per-step loss, per-step parameter gradients and final
    parameters of the fused forward + adjoint kernels vs the reference's own torch-CPU fp64 autograd, <= 1e-8.
    The reference's paraxial code (EPL / EPD, paraxial.py:207-280) stays in the graph on both sides, so the chain
    through the launch plane is part of what is compared."""
    rc.use_torch("cpu", "float64", grad=True)
    want = rc.adam_run(name)
    rc.restore()
    _fused(grad=True)
    got = rc.adam_run(name)
    st = dropin.stats()
    _only_fused(st, grad=True)
    assert st["fused_grad"] >= 3 * 4                      # 3 operands x 4 steps (+ the final evaluation)
    assert np.isfinite(want["grads"]).all() and (want["grads"] != 0).all()
    assert np.max(np.abs(got["losses"] - want["losses"]) / np.abs(want["losses"])) <= 1e-8
    assert np.max(np.abs(got["grads"] - want["grads"]) / np.abs(want["grads"])) <= 1e-8
    assert np.max(np.abs(got["x"] - want["x"]) / np.maximum(np.abs(want["x"]), 1e-12)) <= 1e-8
    assert abs(got["final"] - want["final"]) <= 1e-8 * abs(want["final"])


def _tilt_case(device):
    lens = rc.make_lens("cooke")
    p = torch.nn.Parameter(torch.tensor(0.01, dtype=torch.float64, device=device))
    lens.surface_group.surfaces[3].geometry.cs.rx = p * 1.0
    rays = lens.trace(0.0, 1.0, 0.55, 6, "hexapolar")
    loss = (rays.y[torch.isfinite(rays.y)] ** 2).mean()
    loss.backward()
    return float(loss), float(p.grad)


def test_tilt_variable_goes_through_the_adjoint_kernel():
    """A tilt that depends on an optimisation variable (optimization/variable/tilt.py:59): the general adjoint kernel
    returns d loss / d cos, d sin of the frame's angles and autograd chains them to the angle. Against the reference's
    own op-by-op graph on the same device (stock method), <= 1e-8."""
    rc.use_torch("cuda", "float64", grad=True)
    want = _tilt_case("cuda")
    _fused(grad=True)
    got = _tilt_case("cuda")
    st = dropin.stats()
    assert st.get("fused_grad", 0) == 1 and not any(k.startswith("stock:grad") for k in st), st
    assert abs(got[0] - want[0]) <= 1e-10 * abs(want[0])
    assert want[1] != 0.0 and abs(got[1] - want[1]) <= 1e-8 * abs(want[1]), (got, want)


def _three_mirror_case(device):
    """BASELINE C3's three-mirror system built on the reference's Optic; leaves: the rx tilt of every mirror, the dy
    decenter of the second, the 4 x 4 polynomial coefficients, radius and conic of the first."""
    from optiland import optic as ref_optic
    from optiland_pr_b200 import samples as my_samples
    lens = my_samples.build_three_mirror(ref_optic.Optic(name="TMA"), to_array=lambda a: [list(r) for r in a])
    sf = lens.surface_group.surfaces
    mk = lambda v: torch.tensor(float(v), dtype=torch.float64, device=device, requires_grad=True)
    leaves = []
    for k in (1, 2, 3):
        t = mk(sf[k].geometry.cs.rx)
        sf[k].geometry.cs.rx = t
        leaves.append(t)
    t = mk(sf[2].geometry.cs.y)
    sf[2].geometry.cs.y = t
    leaves.append(t)
    c2 = torch.tensor(rc.to_np(sf[1].geometry.coefficients), dtype=torch.float64, device=device, requires_grad=True)
    sf[1].geometry.coefficients = c2
    leaves.append(c2)
    for nm in ("radius", "k"):
        t = mk(getattr(sf[1].geometry, nm))
        setattr(sf[1].geometry, nm, t)
        leaves.append(t)
    loss = 0.0
    for Hy in (0.0, 1.0, -1.0):
        rays = lens.trace(0.0, Hy, 0.587, 8, "hexapolar")
        loss = loss + ((rays.x - rays.x.mean()) ** 2 + (rays.y - rays.y.mean()) ** 2).mean()
    loss.backward()
    return float(loss), np.concatenate([t.grad.detach().cpu().numpy().reshape(-1) for t in leaves])


def test_three_mirror_tilts_decenters_and_polynomial_coefficients_through_the_adjoint_kernel():
    """tilt, decenter, radius, conic and polynomial-coefficient gradients of the off-axis three-mirror
    freeform system (optimization/variable/tilt.py:59, decenter.py, polynomial_coeff.py:72) from the fused forward +
    general adjoint kernel, against the reference's own torch graph on the same device: <= 1e-8 of the largest entry per
    parameter kind."""
    rc.use_torch("cuda", "float64", grad=True)
    want = _three_mirror_case("cuda")
    _fused(grad=True)
    got = _three_mirror_case("cuda")
    st = dropin.stats()
    assert st.get("fused_grad", 0) == 3 and not any(k.startswith("stock:grad") for k in st), st
    assert abs(got[0] - want[0]) <= 1e-10 * abs(want[0])
    g, w = got[1], want[1]
    assert np.isfinite(w).all() and (w != 0).sum() >= 20
    for sl in (slice(0, 3), slice(3, 4), slice(4, 20), slice(20, 22)):      # tilts, decenter, coefficients, R and k
        assert np.max(np.abs(g[sl] - w[sl])) <= 1e-8 * np.abs(w[sl]).max(), (sl, g[sl], w[sl])


def test_gradient_wanted_for_an_undifferentiated_quantity_uses_the_stock_graph():
    """The normalisation radius of a Zernike surface as an optimisation variable (optimization/variable/
    norm_radius.py): the adjoint kernels do not differentiate it, so the patched method must hand the trace to the
    stock loop (autograd through the reference's own ops), not drop the gradient."""
    from optiland import optic as ref_optic
    import optiland.physical_apertures as ref_ap
    from optiland_pr_b200 import samples as my_samples
    _fused(grad=False)
    lens = my_samples.build_freeform_plate(ref_optic.Optic(name="plate"), ref_ap, "fringe")
    p = torch.nn.Parameter(torch.tensor(9.5, dtype=torch.float64, device="cuda"))
    lens.surface_group.surfaces[1].geometry.norm_radius = p * 1.0
    rays = lens.trace(0.0, 0.0, 0.55, 32, "random")
    ok = torch.isfinite(rays.y)
    loss = (rays.y[ok] ** 2).mean()
    loss.backward()
    st = dropin.stats()
    assert st.get("stock:grad of norm_radius", 0) == 1 and "fused_grad" not in st, st
    assert p.grad is not None and np.isfinite(float(p.grad)) and float(p.grad) != 0.0


def _plate_case(device):
    """The Zernike (noll) + odd-asphere plate on the reference's Optic; leaves: the 12 Zernike coefficients
    (optimization/variable/zernike_coeff.py:71), one odd-asphere coefficient, a tilt and the conic of the Zernike
    surface. Grad MODE stays off: with it on, the reference's own Zernike radial term floor-divides requires-grad
    tensors ((n + m) // 2, zernike/base.py) and torch refuses to differentiate its graph."""
    from optiland import optic as ref_optic
    import optiland.physical_apertures as ref_ap
    from optiland_pr_b200 import samples as my_samples
    lens = my_samples.build_freeform_plate(ref_optic.Optic(name="plate"), ref_ap, "noll")
    sf = lens.surface_group.surfaces
    mk = lambda v: torch.tensor(float(v), dtype=torch.float64, device=device, requires_grad=True)
    cz = torch.tensor(rc.to_np(sf[1].geometry.zernike.coeffs), dtype=torch.float64, device=device, requires_grad=True)
    sf[1].geometry.zernike.coeffs = cz
    co = list(sf[2].geometry.coefficients)
    co[1] = mk(co[1])
    sf[2].geometry.coefficients = co
    rx = mk(sf[2].geometry.cs.rx)
    sf[2].geometry.cs.rx = rx
    k1 = mk(sf[1].geometry.k)
    sf[1].geometry.k = k1
    rays = lens.trace(0.0, 0.0, 0.55, 48, "random")
    ok = torch.isfinite(rays.x) & (rays.i > 0)
    loss = (rays.x[ok] ** 2 + rays.y[ok] ** 2).mean() + 1e-3 * rays.opd[ok].mean()
    loss.backward()
    return float(loss), np.concatenate([t.grad.detach().cpu().numpy().reshape(-1) for t in (cz, co[1], rx, k1)])


def test_zernike_and_odd_asphere_coefficients_through_the_adjoint_kernel():
    """VERDICT r1 item 7, last part: Zernike-coefficient (and odd-asphere, tilt, conic) gradients of the freeform plate
    from the fused forward + general adjoint kernel against the reference's own torch graph on the same device."""
    rc.use_torch("cuda", "float64", grad=False)
    want = _plate_case("cuda")
    _fused(grad=False)
    got = _plate_case("cuda")
    st = dropin.stats()
    assert st.get("fused_grad", 0) == 1 and not any(k.startswith("stock:grad") for k in st), st
    assert abs(got[0] - want[0]) <= 1e-10 * abs(want[0])
    g, w = got[1], want[1]
    assert np.isfinite(w).all() and (w != 0).sum() >= 10
    for sl in (slice(0, 12), slice(12, 13), slice(13, 14), slice(14, 15)):   # Zernike, odd asphere, tilt, conic
        assert np.max(np.abs(g[sl] - w[sl])) <= 1e-8 * np.abs(w[sl]).max(), (sl, g[sl], w[sl])
