"""Value-for-value comparison with the oracle at 2 M rays per BASELINE lens: the NumPy oracle on a
process pool (oracle/pool.py, batch-global Newton counts) against the CUDA kernels through the C ABI. Tolerance mode:
<= 1e-10 relative with identical NaN patterns and vignetting masks (BASELINE.json north_star); strict mode: the
closed-form lens bit for bit (all but the exp() in its intensity), the others <= 1e-12 (their transcendental functions differ by libm ulps)."""
import numpy as np
import pytest
import torch

from oracle import pool, raygen_np
from optiland_pr_b200 import backend as be
from optiland_pr_b200 import samples
from optiland_pr_b200.prescription import pack_surface_group
from optiland_pr_b200.trace import fused_trace

pytestmark = pytest.mark.gpu
N_RAYS = 2_097_152

LENSES = {
    # BASELINE configs: C1 / C4 (Cooke triplet, 3 fields x 3 wavelengths), C2 / C5 (even-asphere double Gauss),
    # C3 (off-axis three-mirror freeform system; Zernike freeform plate with TIR and aperture programs)
    "c1_cooke_triplet": (samples.CookeTriplet, None),
    "c2_even_asphere_double_gauss": (samples.EvenAsphereDoubleGauss, [0.5876]),
    "c3_three_mirror": (samples.ThreeMirrorAnastigmat, None),
    "c3_freeform_plate": (samples.FreeformPlate, None),
}


def _batch(lens, wls, n, seed):
    """n rays: fields x wavelengths x random pupil, through the lens's own ray generator (device)."""
    fields = lens.get_field_coords()
    rng = np.random.default_rng(seed)
    px, py = raygen_np.random_pupil(n, seed=seed)
    f_idx = rng.integers(0, len(fields), n)
    Hx = np.array([f[0] for f in fields])[f_idx]
    Hy = np.array([f[1] for f in fields])[f_idx]
    wl = np.asarray(wls)[rng.integers(0, len(wls), n)]
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    rays = lens.generate_rays(t(Hx), t(Hy), t(px), t(py), t(wl))
    return [c.contiguous() for c in rays.components()]


@pytest.mark.parametrize("name", list(LENSES))
def test_two_million_rays_value_for_value(name):
    be.set_device("cuda:0")
    be.set_precision("float64")
    ctor, wls = LENSES[name]
    lens = ctor()
    wls = [float(w) for w in (wls or lens.wavelengths)]
    ins = _batch(lens, wls, N_RAYS, seed=11)
    packed = pack_surface_group(lens.surface_group.surfaces, wls, record="none")
    rays_np = np.stack([c.cpu().numpy() for c in ins])
    want, k_batch = pool.trace_batch(packed, rays_np)
    nan_w = np.isnan(want[0])
    for fast in (True, False):
        res = fused_trace(packed, ins, fast=fast)
        got = res.out.cpu().numpy()
        assert list(res.nr_iters) == list(k_batch), (res.nr_iters, k_batch)
        assert np.array_equal(np.isnan(got[0]), nan_w), "NaN pattern"
        assert np.array_equal(got[6] == 0, want[6] == 0), "vignetting mask"
        ok = ~nan_w
        worst = 0.0
        for q in range(8):
            scale = np.maximum(np.abs(want[q][ok]), 1.0)
            worst = max(worst, float(np.max(np.abs(got[q][ok] - want[q][ok]) / scale)))
        bar = 1e-10 if fast else 1e-12
        assert worst <= bar, (name, fast, worst)
        if not fast and name == "c1_cooke_triplet":
            # positions, directions and OPD use +, -, *, /, sqrt only: the strict kernel is bit-identical to the oracle;
            # the intensity goes through exp() (Beer-Lambert in the absorbing glasses): libm against CUDA, <= 2 ulp
            for q in (0, 1, 2, 3, 4, 5, 7):
                assert np.array_equal(got[q][ok], want[q][ok]), q
            assert np.abs(got[6][ok] - want[6][ok]).max() <= 4.5e-16
    assert (~nan_w).sum() > 0.8 * N_RAYS
