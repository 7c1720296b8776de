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
