"""GPU: spot.SpotDiagram (all fields x wavelengths in one fused launch + trace + spot kernel per pass)
against the reference's SpotDiagram numbers (tests/golden/spot_diagram.npz, oracle/make_golden_spot.py), in tolerance
mode (fused, per-group Newton counts) and in strict mode (per-pair traces + the two-pass spot kernel)."""
import numpy as np
import pytest

from helpers import load_golden
from optiland_pr_b200 import samples, spot

pytestmark = pytest.mark.gpu


def _numbers(sd):
    c = np.array([[float(a), float(b)] for a, b in sd.centroid()])
    r = np.array([[float(v) for v in row] for row in sd.rms_spot_radius()])
    g = np.array([[float(v) for v in row] for row in sd.geometric_spot_radius()])
    return c, r, g


@pytest.mark.parametrize("name,make", [("cooke", samples.CookeTriplet), ("eadg", samples.EvenAsphereDoubleGauss)])
@pytest.mark.parametrize("fast", [True, False])
def test_spot_diagram_matches_reference(name, make, fast):
    from optiland_pr_b200 import backend as be
    be.set_precision("float64")
    g = load_golden("spot_diagram")
    lens = make()
    assert np.array_equal(np.array(lens.wavelengths), g[f"{name}_wavelengths"])
    lens.surface_group.fast_math = fast
    sd = spot.SpotDiagram(lens, num_rings=6, distribution="hexapolar")
    assert sd.path == ("fused" if fast else "per_pair")
    c, r, geo = _numbers(sd)
    # the on-axis centroid is a sum of +-1e-16 terms: absolute tolerance on the centroid, relative on the radii
    assert np.allclose(c, g[f"{name}_centroid"], rtol=0, atol=1e-10)
    assert np.allclose(r, g[f"{name}_rms"], rtol=1e-9, atol=0)
    assert np.allclose(geo, g[f"{name}_geo"], rtol=1e-9, atol=0)


def test_spot_diagram_subset_and_device_grid():
    """Chosen fields / wavelengths and a caller's pupil grid: same numbers as tracing each pair on its own."""
    import torch

    from optiland_pr_b200 import backend as be
    be.set_precision("float64")
    lens = samples.EvenAsphereDoubleGauss()
    rng = np.random.default_rng(3)
    r_, th = rng.uniform(size=5000), rng.uniform(0, 2 * np.pi, size=5000)
    px, py = np.sqrt(r_) * np.cos(th), np.sqrt(r_) * np.sin(th)
    fields = [(0.0, 0.3), (0.0, 1.0)]
    wls = [0.6563, 0.5876]
    sd = spot.SpotDiagram(lens, fields=fields, wavelengths=wls, distribution=(px, py))
    assert sd.path == "fused" and sd.ref_index == 1
    _, r, geo = _numbers(sd)
    lens.surface_group.fast_math = False
    Px, Py = torch.tensor(px, device="cuda"), torch.tensor(py, device="cuda")
    for f, (hx, hy) in enumerate(fields):
        ref = lens.trace(np.array([hx]), np.array([hy]), 0.5876, distribution=(Px, Py))
        cen = spot.spot_radii(ref.x, ref.y, ref.i)
        c = torch.stack([cen["cx"], cen["cy"]], dim=1)
        for w, wl in enumerate(wls):
            rr = lens.trace(np.array([hx]), np.array([hy]), wl, distribution=(Px, Py))
            want = spot.spot_radii(rr.x, rr.y, rr.i, reference_centroid=c)
            assert abs(r[f, w] - float(want["rms"])) <= 1e-9 * float(want["rms"])
            assert abs(geo[f, w] - float(want["geo"])) <= 1e-9 * float(want["geo"])
