"""Ray launch: the NumPy restatement (CPU) and the CUDA launch kernel (GPU) against rays produced by the
reference's own RayGenerator (stored in tests/golden/cooke_triplet.npz: 3 fields x 3 wavelengths x 150
RandomDistribution(seed=0) pupil points)."""
import numpy as np
import pytest

from helpers import load_golden
from oracle import raygen_np
from optiland_pr_b200 import samples


def _blocks():
    g = load_golden("cooke_triplet")
    lens = samples.CookeTriplet()
    px, py = raygen_np.random_pupil(150, seed=0)
    return g, lens, px, py


def test_numpy_ray_launch_matches_reference_bits():
    g, lens, px, py = _blocks()
    surf_z = [float(s.geometry.cs.z) for s in lens.surface_group.surfaces[1:]]
    k = 0
    for (hx, hy) in lens.get_field_coords():
        for w in lens.wavelengths:
            got = np.stack(raygen_np.generate_rays(hx, hy, px, py, w, EPL=g["paraxial"][0], EPD=g["paraxial"][1],
                                                   max_field=lens.max_field, surface_z=surf_z))
            assert np.array_equal(got, g["rays_in"][:, k * 150:(k + 1) * 150])
            k += 1


@pytest.mark.gpu
def test_cuda_ray_launch_matches_reference_bits():
    import torch

    g, lens, px, py = _blocks()
    fields = lens.get_field_coords()
    Px = torch.tensor(px, device="cuda")
    Py = torch.tensor(py, device="cuda")
    for wi, w in enumerate(lens.wavelengths):
        rays = lens.launch_rays(np.array([f[0] for f in fields]), np.array([f[1] for f in fields]), Px, Py, w)
        got = torch.stack(rays.components()).cpu().numpy()
        for fi in range(len(fields)):
            k = fi * len(lens.wavelengths) + wi
            assert np.array_equal(got[:, fi * 150:(fi + 1) * 150], g["rays_in"][:, k * 150:(k + 1) * 150]), (fi, wi)


@pytest.mark.gpu
def test_pipelined_host_pupil_equals_resident_trace():
    """Optic.trace with a pinned HOST pupil grid (chunked copy overlapped with launch + trace) gives the
    same bits as the same call with the grid already on the device."""
    import torch

    from optiland_pr_b200 import backend as be

    be.set_precision("float64")
    lens = samples.EvenAsphereDoubleGauss()
    lens.surface_group.record = "last"
    lens.pipeline_min_rays = 1 << 16
    px, py = raygen_np.random_pupil(300_000, seed=5)
    Hx, Hy = [0.0, 0.0, 0.0], [0.0, 0.7, 1.0]
    a = lens.trace(Hx, Hy, 0.5876, distribution=(torch.from_numpy(px).pin_memory(), torch.from_numpy(py).pin_memory()))
    xa = torch.stack([a.x, a.y, a.z, a.L, a.M, a.N, a.i, a.opd, a.L0, a.M0, a.N0]).cpu().numpy()
    last_a = lens.surface_group.x[-1].cpu().numpy()
    b = lens.trace(Hx, Hy, 0.5876, distribution=(torch.from_numpy(px).cuda(), torch.from_numpy(py).cuda()))
    xb = torch.stack([b.x, b.y, b.z, b.L, b.M, b.N, b.i, b.opd, b.L0, b.M0, b.N0]).cpu().numpy()
    assert np.array_equal(xa, xb, equal_nan=True)
    assert np.array_equal(last_a, lens.surface_group.x[-1].cpu().numpy())


# ---- launch cases beyond the infinite-conjugate, unvignetted one (tests/golden/raygen_cases.npz, generated from the
# ---- unmodified reference by oracle/make_golden_raygen.py) -----------------------------------------------------------
WL = 0.5876


def _case_lens(name):
    from optiland_pr_b200.optic import Optic
    if name == "vig":
        ref = samples.CookeTriplet()
        lens = Optic()
        lens.surface_group = ref.surface_group
        lens.set_aperture("EPD", 10)
        lens.set_field_type("angle")
        lens.add_field(y=0)
        lens.add_field(y=14, vx=0.1, vy=0.25)
        lens.add_field(y=16, x=12, vx=0.2, vy=0.4)
    else:
        lens = Optic()
        lens.add_surface(index=0, radius=np.inf if name == "finite" else -400.0, thickness=120.0)
        lens.add_surface(index=1, radius=45.0, thickness=6.0, material="N-BK7", is_stop=True)
        lens.add_surface(index=2, radius=-60.0, conic=-0.8, thickness=70.0)
        lens.add_surface(index=3)
        lens.set_aperture("EPD", 12.0)
        if name == "finite":
            lens.set_field_type("angle")
            for x, y in ((0, 0), (0, 3.0), (2.0, 4.0)):
                lens.add_field(y=y, x=x)
        else:
            lens.set_field_type("object_height")
            for x, y in ((0, 0), (0, 5.0), (-3.0, 7.0)):
                lens.add_field(y=y, x=x, vx=0.05 if y else 0.0, vy=0.1 if y else 0.0)
    lens.add_wavelength(WL, is_primary=True)
    return lens


CASES = ("vig", "finite", "height")


@pytest.mark.parametrize("name", CASES)
def test_numpy_ray_launch_cases_match_reference_bits(name):
    g = load_golden("raygen_cases")
    lens = _case_lens(name)
    assert lens.max_field == float(g[f"{name}_max_field"])
    coords = np.array(lens.get_field_coords())
    assert np.array_equal(coords, g[f"{name}_coords"])
    surf_z = [float(s.geometry.cs.z) for s in lens.surface_group.surfaces[1:]]
    obj = lens.surface_group.surfaces[0]
    for f, (hx, hy) in enumerate(coords):
        vxf, vyf = raygen_np.vig_factor(lens.fields, lens.vignetting, lens.max_field, hx, hy)
        got = np.stack(raygen_np.generate_rays(
            hx, hy, g["px"], g["py"], WL, EPL=float(g[f"{name}_EPL"]), EPD=float(g[f"{name}_EPD"]),
            max_field=lens.max_field, surface_z=surf_z, vx=1 - vxf, vy=1 - vyf,
            object_z=None if obj.is_infinite else float(obj.geometry.cs.z),
            field_type=lens.field_type, object_radius=float(getattr(obj.geometry, "radius", np.inf)),
            object_conic=float(getattr(obj.geometry, "k", 0.0))))
        assert np.array_equal(got, g[f"{name}_rays_in"][f]), (name, f)


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_cuda_ray_launch_cases_match_reference_bits(name):
    """olrt_generate_rays_* (vignetting factors, finite object, object heights) against the reference's RayGenerator,
    and the torch expression path of Optic.generate_rays (trace_generic) against the same rays."""
    import torch

    from optiland_pr_b200 import backend as be
    be.set_precision("float64")
    g = load_golden("raygen_cases")
    lens = _case_lens(name)
    lens.set_paraxial(float(g[f"{name}_EPL"]), float(g[f"{name}_EPD"]))
    coords = np.array(lens.get_field_coords())
    Px, Py = torch.tensor(g["px"], device="cuda"), torch.tensor(g["py"], device="cuda")
    rays = lens.launch_rays(coords[:, 0], coords[:, 1], Px, Py, WL)
    got = torch.stack(rays.components()).cpu().numpy().reshape(8, len(coords), -1).transpose(1, 0, 2)
    assert np.array_equal(got, g[f"{name}_rays_in"])
    for f, (hx, hy) in enumerate(coords):
        r2 = lens.generate_rays(hx, hy, g["px"], g["py"], WL)
        got2 = torch.stack(r2.components()).cpu().numpy()
        assert np.allclose(got2, g[f"{name}_rays_in"][f], rtol=1e-14, atol=1e-14), (name, f)


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
@pytest.mark.parametrize("fast", [True, False])
def test_traced_launch_cases_match_reference(name, fast):
    """Optic.trace end to end for the three launch cases: strict mode (launch kernel + strict trace kernel) gives the
    reference's bits, tolerance mode (ONE fused launch + trace kernel over all fields) <= 1e-10 with the same masks."""
    import torch

    from optiland_pr_b200 import backend as be
    be.set_precision("float64")
    g = load_golden("raygen_cases")
    lens = _case_lens(name)
    lens.set_paraxial(float(g[f"{name}_EPL"]), float(g[f"{name}_EPD"]))
    lens.surface_group.fast_math = fast
    lens.surface_group.record = "last"
    coords = np.array(lens.get_field_coords())
    Px, Py = torch.tensor(g["px"], device="cuda"), torch.tensor(g["py"], device="cuda")
    rays = lens.trace(coords[:, 0], coords[:, 1], WL, distribution=(Px, Py))
    got = torch.stack([rays.x, rays.y, rays.z, rays.L, rays.M, rays.N, rays.i, rays.opd]).cpu().numpy()
    want = g[f"{name}_rays_out"].transpose(1, 0, 2).reshape(8, -1)
    assert np.array_equal(np.isnan(got), np.isnan(want))
    if fast:
        m = ~np.isnan(want)
        assert np.max(np.abs(got[m] - want[m]) / np.maximum(np.abs(want[m]), 1.0)) <= 1e-10
        assert np.array_equal(got[6] == 0, want[6] == 0)
    else:      # + - * / sqrt bit for bit; the intensity goes through exp (absorbing glass): 1e-14
        for q in (0, 1, 2, 3, 4, 5, 7):
            assert np.array_equal(got[q], want[q], equal_nan=True), q
        assert np.allclose(got[6], want[6], rtol=1e-14, atol=0, equal_nan=True)
