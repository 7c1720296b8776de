"""GPU: chief-ray wavefront / OPD maps (optiland_pr_b200/wavefront.py, olrt_opd_epilogue_*) against the output of the
unmodified reference's Wavefront(..., strategy="chief_ray").get_data (tests/golden/wavefront_cooke.npz)."""
import numpy as np
import pytest
import torch

from helpers import load_golden
from optiland_pr_b200 import _lib as L
from optiland_pr_b200 import backend as be
from optiland_pr_b200 import samples
from optiland_pr_b200.rays import RealRays
from optiland_pr_b200.wavefront import Wavefront, opd_epilogue

pytestmark = pytest.mark.gpu
NAMES = ("x", "y", "z", "L", "M", "N", "opd")


def test_epilogue_kernel_on_the_reference_rays_is_bit_exact():
    """Same traced rays in -> the reference's bits out (fp64, strict build, IEEE division and sqrt)."""
    from oracle import wavefront_np
    g = load_golden("wavefront_cooke")
    for k, f in enumerate(g["fields"]):
        r = g[f"f{k}_rays"]
        t = lambda a: torch.tensor(np.ascontiguousarray(a), device="cuda")
        rays = RealRays(t(r[0]), t(r[1]), t(r[2]), t(r[3]), t(r[4]), t(r[5]), t(r[7]), torch.full((r.shape[1],), 0.55, device="cuda"))
        rays.opd = t(r[6])
        c = g[f"f{k}_chief"][:, 0]
        R = float(g[f"f{k}_radius"])
        ux, uy = wavefront_np.tilt_direction(f[0], f[1], float(g["max_field"]))
        sph = L.OlrtRefSphere(c[0], c[1], c[2], R, float(g["n_image"]), 0.0, 0.55, float(ux), float(uy), float(g["EPD"]))
        ch = RealRays(*[t(c[q:q + 1]) for q in range(6)], torch.ones(1, device="cuda"), torch.full((1,), 0.55, device="cuda"))
        ch.opd = t(c[6:7])
        opd_ref, _ = opd_epilogue(ch, None, None, sph, raw=True, want_pupil=False)
        sph.opd_ref = float(opd_ref[0])
        opd, pupil = opd_epilogue(rays, t(g["px"]), t(g["py"]), sph)
        assert np.array_equal(opd.cpu().numpy(), g[f"f{k}_opd"], equal_nan=True)
        assert np.array_equal(pupil.cpu().numpy(), g[f"f{k}_pupil"], equal_nan=True)


def test_wavefront_through_the_public_api():
    """Trace (tolerance mode) + epilogue: OPD map within 1e-8 waves of the reference, pupil points within 1e-10."""
    be.set_precision("float64")
    g = load_golden("wavefront_cooke")
    lens = samples.CookeTriplet()
    lens.set_paraxial(EPL=float(g["EPL"]), EPD=float(g["EPD"]), XPL=float(g["XPL"]))
    wf = Wavefront(lens, [tuple(f) for f in g["fields"]], [0.55], distribution=(g["px"], g["py"]))
    for k, f in enumerate(g["fields"]):
        d = wf.get_data(tuple(f), 0.55)
        assert abs(d.radius - float(g[f"f{k}_radius"])) <= 1e-10 * d.radius
        assert np.max(np.abs(d.opd.cpu().numpy() - g[f"f{k}_opd"])) <= 1e-8
        got = torch.stack([d.pupil_x, d.pupil_y, d.pupil_z]).cpu().numpy()
        assert np.max(np.abs(got - g[f"f{k}_pupil"]) / np.maximum(np.abs(g[f"f{k}_pupil"]), 1.0)) <= 1e-10
        assert np.array_equal(d.intensity.cpu().numpy() == 0, g[f"f{k}_intensity"] == 0)
