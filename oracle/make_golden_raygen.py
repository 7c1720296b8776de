"""TEST INFRASTRUCTURE ONLY -- tests/golden/raygen_cases.npz from the UNMODIFIED reference: rays launched by
RayGenerator.generate_rays (rays/ray_generator.py:28-106) and traced by SurfaceGroup.trace for the launch cases the
Cooke-triplet fixture does not hold:
  vig     Cooke triplet, object at infinity, vignetting factors on the off-axis fields, one field off the y axis
  finite  a finite-conjugate singlet, angle fields (AngleField.get_ray_origins, field_types.py:158-168)
  height  the same singlet with a CURVED object surface and object-height fields (field_types.py:245-272)
Each case stores the paraxial numbers, the pupil grid, the launched rays [F, 8, P] and the ray state at the image
[F, 8, P] (x y z L M N intensity opd).     python oracle/make_golden_raygen.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_shim  # noqa: E402

ref_shim.install()
import optiland.backend as be  # noqa: E402
from optiland import optic  # noqa: E402
from optiland.rays import RayGenerator  # noqa: E402
from optiland.samples.objectives import CookeTriplet  # noqa: E402

IN = ("x", "y", "z", "L", "M", "N", "i", "w")
OUT = ("x", "y", "z", "L", "M", "N", "i", "opd")
WL = 0.5876


def singlet(field_type, obj_radius):
    lens = optic.Optic()
    lens.add_surface(index=0, radius=obj_radius, thickness=120.0)
    lens.add_surface(index=1, radius=45.0, thickness=6.0, material="N-BK7", is_stop=True)
    lens.add_surface(index=2, radius=-60.0, conic=-0.8, thickness=70.0)
    lens.add_surface(index=3)
    lens.set_aperture(aperture_type="EPD", value=12.0)
    lens.set_field_type(field_type=field_type)
    if field_type == "angle":
        for x, y in ((0, 0), (0, 3.0), (2.0, 4.0)):
            lens.add_field(y=y, x=x)
    else:
        for x, y in ((0, 0), (0, 5.0), (-3.0, 7.0)):
            lens.add_field(y=y, x=x, vx=0.05 if y else 0.0, vy=0.1 if y else 0.0)
    lens.add_wavelength(value=WL, is_primary=True)
    return lens


def vig_cooke():
    ref = CookeTriplet()
    lens = optic.Optic()
    lens.surface_group = ref.surface_group
    lens.set_aperture(aperture_type="EPD", value=10)
    lens.set_field_type(field_type="angle")
    lens.add_field(y=0)
    lens.add_field(y=14, vx=0.1, vy=0.25)
    lens.add_field(y=16, x=12, vx=0.2, vy=0.4)
    lens.add_wavelength(value=WL, is_primary=True)
    return lens


def case(lens, px, py):
    gen_in, traced = [], []
    for hx, hy in lens.fields.get_field_coords():
        rays = RayGenerator(lens).generate_rays(hx, hy, px, py, WL)
        gen_in.append(np.stack([np.asarray(getattr(rays, k), dtype=np.float64) * np.ones_like(px) for k in IN]))
        with np.errstate(all="ignore"):
            lens.surface_group.trace(rays)
        traced.append(np.stack([np.asarray(getattr(rays, k), dtype=np.float64) for k in OUT]))
    return dict(EPL=float(lens.paraxial.EPL()), EPD=float(lens.paraxial.EPD()), max_field=float(lens.fields.max_field),
                coords=np.array(lens.fields.get_field_coords(), dtype=np.float64),
                rays_in=np.stack(gen_in), rays_out=np.stack(traced))


def main():
    be.set_backend("numpy")
    rng = np.random.default_rng(11)
    r_, th = rng.uniform(size=96), rng.uniform(0, 2 * np.pi, size=96)
    px, py = np.sqrt(r_) * np.cos(th), np.sqrt(r_) * np.sin(th)
    px[0] = py[0] = 0.0
    out = dict(px=px, py=py, wavelength=WL)
    for name, lens in (("vig", vig_cooke()), ("finite", singlet("angle", be.inf)),
                       ("height", singlet("object_height", -400.0))):
        for k, v in case(lens, px, py).items():
            out[f"{name}_{k}"] = v
    path = os.path.join(ROOT, "tests", "golden", "raygen_cases.npz")
    np.savez_compressed(path, **out)
    print(path, {k: getattr(v, "shape", v) for k, v in out.items()})


if __name__ == "__main__":
    main()
