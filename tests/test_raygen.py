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
