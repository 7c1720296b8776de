"""TEST INFRASTRUCTURE ONLY -- generate tests/golden/*.npz by running the UNMODIFIED reference.

    python oracle/make_golden.py            # needs /root/reference (build container only)

Each fixture holds, for one lens + ray batch: the input rays exactly as the reference's own
RayGenerator produced them, the reference NumPy-float64 outputs of `surface_group.trace`
(every recorded surface: x y z L M N intensity opd), and the packed prescription of the
reference's objects. The GPU box has no reference; the `-m gpu` tests rebuild the same lens
with this package's mirror classes, trace the stored rays with the CUDA kernel and compare.
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_shim  # noqa: E402

ref_shim.install()

import optiland.backend as be  # noqa: E402
from optiland import optic as ref_optic  # noqa: E402
from optiland.distribution import create_distribution  # noqa: E402
from optiland.samples.objectives import CookeTriplet, DoubleGauss  # noqa: E402
from optiland.samples.simple import AsphericSinglet  # noqa: E402

from optiland_pr_b200 import samples as my_samples  # noqa: E402
from optiland_pr_b200.prescription import pack_surface_group  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")
FIELDS8 = ("x", "y", "z", "L", "M", "N", "intensity", "opd")


def ref_even_asphere_double_gauss():
    """BASELINE C2 lens built with the reference's own add_surface (see samples.py)."""
    lens = ref_optic.Optic()
    rows = [
        dict(index=1, radius=56.20238, thickness=8.75, material="N-SSK2"),
        dict(index=2, radius=152.28580, thickness=0.5),
        dict(index=3, radius=37.68262, thickness=12.5, material="N-SK2"),
        dict(index=4, radius=be.inf, thickness=3.8, material=("F5", "schott")),
        dict(index=5, radius=24.23130, thickness=16.369445),
        dict(index=6, radius=be.inf, thickness=13.747957, is_stop=True),
        dict(index=7, radius=-28.37731, thickness=3.8, material=("F5", "schott")),
        dict(index=8, radius=be.inf, thickness=11, material="N-SK16"),
        dict(index=9, radius=-37.92546, thickness=0.5),
        dict(index=10, radius=177.41176, thickness=7, material="N-SK16"),
        dict(index=11, radius=-79.41143, thickness=61.487536),
    ]
    lens.add_surface(index=0, radius=be.inf, thickness=be.inf)
    for row in rows:
        if row["index"] in my_samples.DOUBLE_GAUSS_ASPHERES:
            row = dict(row, surface_type="even_asphere", conic=0.0,
                       coefficients=list(my_samples.DOUBLE_GAUSS_ASPHERES[row["index"]]))
        lens.add_surface(**row)
    lens.add_surface(index=12)
    lens.set_aperture(aperture_type="imageFNO", value=5)
    lens.set_field_type(field_type="angle")
    for fy in (0, 10, 14):
        lens.add_field(y=fy)
    lens.add_wavelength(value=0.4861)
    lens.add_wavelength(value=0.5876, is_primary=True)
    lens.add_wavelength(value=0.6563)
    return lens


def gen_rays(lens, fields, wavelengths, dist_name, num, seed=None):
    """One batch = every (field, wavelength) pair, pupil from the reference's distribution."""
    dist = create_distribution(dist_name)
    if dist_name == "random":
        dist = type(dist)(seed=seed)
    dist.generate_points(num)
    parts = []
    for (hx, hy) in fields:
        for w in wavelengths:
            n = len(dist.x)
            parts.append(lens.ray_tracer.ray_generator.generate_rays(
                np.full(n, hx), np.full(n, hy), dist.x, dist.y, w))
    cat = lambda name: np.concatenate([np.asarray(getattr(p, name), dtype=np.float64) for p in parts])
    from optiland.rays import RealRays
    return RealRays(cat("x"), cat("y"), cat("z"), cat("L"), cat("M"), cat("N"), cat("i"), cat("w"))


def run_case(name, lens, rays, wavelengths, extra=None):
    rays_in = np.stack([np.asarray(getattr(rays, k), dtype=np.float64).copy()
                        for k in ("x", "y", "z", "L", "M", "N", "i", "w")])
    sg = lens.surface_group
    with np.errstate(all="ignore"):
        sg.trace(rays)
    rec = np.stack([np.stack([np.asarray(getattr(s, k), dtype=np.float64) for k in FIELDS8])
                    for s in sg.surfaces[1:]])
    packed = pack_surface_group(sg.surfaces, wavelengths)
    nr_iters = []  # Newton updates the reference took: sag evaluations - 1 is not observable; use the oracle
    paraxial = np.array([float(lens.paraxial.EPL()), float(lens.paraxial.EPD())])
    out = dict(rays_in=rays_in, records=rec, wavelengths=np.asarray(wavelengths, dtype=np.float64),
               ops=packed.ops, presc=packed.presc, paraxial=paraxial)
    if extra:
        out.update(extra)
    os.makedirs(GOLDEN, exist_ok=True)
    path = os.path.join(GOLDEN, name + ".npz")
    np.savez_compressed(path, **out)
    nan = int(np.isnan(rec[-1][0]).sum())
    print(f"{name}: rays={rays_in.shape[1]} surfaces={rec.shape[0]} nan_rays={nan} "
          f"clipped={(rec[-1][6] == 0).sum()} -> {os.path.getsize(path) / 1024:.0f} KiB")


def main():
    be.set_backend("numpy")
    # C1: Cooke triplet, 3 fields x 3 wavelengths in one batch (W = 3)
    lens = CookeTriplet()
    wl = [0.48, 0.55, 0.65]
    rays = gen_rays(lens, lens.fields.get_field_coords(), wl, "random", 150, seed=0)
    run_case("cooke_triplet", lens, rays, wl)
    # stock double Gauss (Plane + radius-inf surfaces)
    lens = DoubleGauss()
    wl = [0.4861, 0.5876, 0.6563]
    rays = gen_rays(lens, lens.fields.get_field_coords(), wl, "hexapolar", 6)
    run_case("double_gauss", lens, rays, wl)
    # C2: even-asphere double Gauss (Newton-Raphson surfaces)
    lens = ref_even_asphere_double_gauss()
    rays = gen_rays(lens, lens.fields.get_field_coords(), [0.5876], "random", 400, seed=1)
    run_case("even_asphere_double_gauss", lens, rays, [0.5876])
    # aspheric singlet
    lens = AsphericSinglet()
    rays = gen_rays(lens, [(0.0, 0.0)], [0.587], "hexapolar", 12)
    run_case("aspheric_singlet", lens, rays, [0.587])
    # C3: off-axis three-mirror anastigmat (polynomial freeforms, rx tilts, dy decenters, mirrors)
    lens = my_samples.build_three_mirror(ref_optic.Optic(name="TMA"), to_array=np.array)
    wl = [0.486, 0.587, 0.656]
    rays = gen_rays(lens, lens.fields.get_field_coords(), wl, "random", 120, seed=2)
    run_case("three_mirror", lens, rays, wl)
    # Zernike / odd-asphere / tilted / apertured / absorbing plate with TIR, one fixture per index scheme
    import optiland.physical_apertures as ref_ap
    for zt in ("fringe", "noll", "standard"):
        lens = my_samples.build_freeform_plate(ref_optic.Optic(name="plate"), ref_ap, zt)
        rays = gen_rays(lens, lens.fields.get_field_coords(), [0.55], "random", 700, seed=3)
        run_case("freeform_plate_" + zt, lens, rays, [0.55])
    # SURVEY.md 8f N4 geometries: biconic / toroidal (anamorphic system) and a Chebyshev freeform
    lens = my_samples.build_anamorphic(ref_optic.Optic(name="anamorphic"))
    rays = gen_rays(lens, lens.fields.get_field_coords(), [0.55], "random", 300, seed=4)
    run_case("anamorphic", lens, rays, [0.55])
    lens = my_samples.build_chebyshev_plate(ref_optic.Optic(name="chebyshev"), to_array=np.array)
    rays = gen_rays(lens, lens.fields.get_field_coords(), [0.55], "random", 300, seed=5)
    run_case("chebyshev_plate", lens, rays, [0.55])
    # thin lenses + SimpleCoating (interaction models / coatings of 8f N4)
    from optiland.coatings import SimpleCoating
    lens = my_samples.build_coated_relay(ref_optic.Optic(name="coated_relay"), SimpleCoating)
    rays = gen_rays(lens, lens.fields.get_field_coords(), [0.55], "random", 300, seed=6)
    run_case("coated_relay", lens, rays, [0.55])


if __name__ == "__main__":
    main()
