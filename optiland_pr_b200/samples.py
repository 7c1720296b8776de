"""Sample lenses used by the BASELINE configs, built on this package's Optic.

Prescription DATA follows the reference's samples (optiland/samples/objectives.py:46-114,
optiland/samples/simple.py:161-190). EPL / EPD are the reference's paraxial values
(optiland/paraxial.py:207-280) evaluated once with oracle/make_golden.py and recorded here,
because the paraxial tracer is outside the hot path.

`EvenAsphereDoubleGauss` is BASELINE config C2/C5: the double Gauss with surfaces 1, 5, 7, 11
switched to `even_asphere` (three coefficients each), as SURVEY.md 8(a) proposes; it does not
exist in the reference, so oracle/make_golden.py builds the identical lens with the
reference's own `Optic.add_surface` to produce golden vectors.
"""
from __future__ import annotations

import math

from .optic import Optic

INF = math.inf


class CookeTriplet(Optic):
    def __init__(self):
        super().__init__("CookeTriplet")
        self.add_surface(index=0, radius=INF, thickness=INF)
        self.add_surface(index=1, radius=22.01359, thickness=3.25896, material="SK16")
        self.add_surface(index=2, radius=-435.76044, thickness=6.00755)
        self.add_surface(index=3, radius=-22.21328, thickness=0.99997, material=("F2", "schott"))
        self.add_surface(index=4, radius=20.29192, thickness=4.75041, is_stop=True)
        self.add_surface(index=5, radius=79.68360, thickness=2.95208, material="SK16")
        self.add_surface(index=6, radius=-18.39533, thickness=42.20778)
        self.add_surface(index=7)
        self.set_aperture("EPD", 10)
        self.set_field_type("angle")
        for fy in (0, 14, 20):
            self.add_field(y=fy)
        self.add_wavelength(0.48)
        self.add_wavelength(0.55, is_primary=True)
        self.add_wavelength(0.65)
        self.set_paraxial(EPL=11.512158673746795, EPD=10.0)


DOUBLE_GAUSS_ASPHERES = {
    # surface index -> even-asphere coefficients (r^2, r^4, r^6), sized like AsphericSinglet's
    1: [0.0, -2.0e-7, 1.5e-10],
    5: [0.0, 1.2e-6, -3.0e-9],
    7: [0.0, -1.5e-6, 2.5e-9],
    11: [0.0, 1.0e-7, -0.5e-10],
}


def _double_gauss(optic: Optic, aspheres):
    rows = [
        dict(index=1, radius=56.20238, thickness=8.75, material="N-SSK2"),
        dict(index=2, radius=152.28580, thickness=0.5),
        dict(index=3, radius=37.68262, thickness=12.5, material="N-SK2"),
        dict(index=4, radius=INF, thickness=3.8, material=("F5", "schott")),
        dict(index=5, radius=24.23130, thickness=16.369445),
        dict(index=6, radius=INF, thickness=13.747957, is_stop=True),
        dict(index=7, radius=-28.37731, thickness=3.8, material=("F5", "schott")),
        dict(index=8, radius=INF, thickness=11, material="N-SK16"),
        dict(index=9, radius=-37.92546, thickness=0.5),
        dict(index=10, radius=177.41176, thickness=7, material="N-SK16"),
        dict(index=11, radius=-79.41143, thickness=61.487536),
    ]
    optic.add_surface(index=0, radius=INF, thickness=INF)
    for row in rows:
        if row["index"] in aspheres:
            row = dict(row, surface_type="even_asphere", conic=0.0, coefficients=aspheres[row["index"]])
        optic.add_surface(**row)
    optic.add_surface(index=12)
    optic.set_aperture("imageFNO", 5)
    optic.set_field_type("angle")
    for fy in (0, 10, 14):
        optic.add_field(y=fy)
    optic.add_wavelength(0.4861)
    optic.add_wavelength(0.5876, is_primary=True)
    optic.add_wavelength(0.6563)
    optic.set_paraxial(EPL=57.911999997719626, EPD=20.00074957416745)


class DoubleGauss(Optic):
    def __init__(self):
        super().__init__("DoubleGauss")
        _double_gauss(self, {})


class EvenAsphereDoubleGauss(Optic):
    def __init__(self):
        super().__init__("EvenAsphereDoubleGauss")
        _double_gauss(self, DOUBLE_GAUSS_ASPHERES)


class AsphericSinglet(Optic):
    def __init__(self):
        super().__init__("AsphericSinglet")
        self.add_surface(index=0, radius=INF, thickness=INF)
        self.add_surface(index=1, thickness=7, radius=20.0, is_stop=True, material="N-SF11",
                         surface_type="even_asphere", conic=0.0,
                         coefficients=[-2.248851e-4, -4.690412e-6, -6.404376e-8])
        self.add_surface(index=2, thickness=21.56201105)
        self.add_surface(index=3)
        self.set_aperture("EPD", 20.0)
        self.set_field_type("angle")
        self.add_field(y=0)
        self.add_wavelength(0.587, is_primary=True)
        self.set_paraxial(EPL=0.0, EPD=20.0)


class TiltedSinglet(AsphericSinglet):
    """The aspheric singlet with both surfaces decentred and tilted (oracle/make_golden_grads.py: ref_tilted_singlet):
    refraction through tilted frames, the tilt variables of optimization/variable/tilt.py."""

    def __init__(self):
        super().__init__()
        s1, s2 = self.surface_group.surfaces[1], self.surface_group.surfaces[2]
        s1.geometry.cs.rx, s1.geometry.cs.ry, s1.geometry.cs.rz = 0.03, -0.02, 0.4
        s1.geometry.cs.x, s1.geometry.cs.y = 0.15, -0.1
        s2.geometry.cs.rx, s2.geometry.cs.ry = -0.025, 0.015
        s2.geometry.cs.y = 0.2
        self.surface_group.invalidate()


# ---- builders shared with oracle/make_golden.py: the SAME calls build the lens on the reference's Optic and
# ---- on this package's Optic (`lens` is either; `ap` is the module holding the aperture classes)
TMA_COEFFS = (
    [[2.05790534e-05, -2.04095791e-05, 4.55045366e-05, 1.47183545e-05],
     [3.64010850e-08, -9.18266382e-06, 3.12457789e-07, -3.43010757e-08],
     [-1.18212287e-04, 2.97963986e-06, -1.31354947e-06, 7.34797747e-07],
     [-1.09979245e-06, 8.55358588e-07, -3.94583506e-07, -5.01115270e-08]],
    [[-2.05802662e-05, 2.05376130e-05, -2.91883756e-05, 4.01274602e-05],
     [-2.94473576e-08, -1.63268440e-06, -5.35655902e-06, -1.50920309e-05],
     [4.87006021e-05, -1.70308923e-05, 2.12362562e-05, 1.51092539e-05],
     [-1.04657810e-05, -2.25376153e-05, -8.02314506e-06, 3.02373985e-06]],
    [[2.05787912e-05, -2.05604065e-05, 2.54616632e-05, -3.62841214e-05],
     [6.24577550e-10, 1.09936457e-07, -2.69199040e-06, -1.05068667e-05],
     [-3.28005504e-05, 1.88336844e-05, -3.34058539e-05, 4.93261709e-06],
     [6.09659324e-07, -1.58589636e-06, -1.16163570e-05, -2.26243673e-05]],
)


def build_three_mirror(lens, to_array=lambda a: a):
    """Off-axis three-mirror anastigmat of docs/gallery/reflective/three_mirror_anastigmat.ipynb: polynomial
    4x4 freeform mirrors, rx tilts, dy decenters (BASELINE config C3)."""
    lens.set_aperture(aperture_type="EPD", value=10)
    lens.set_field_type(field_type="angle")
    for fy in (0, 1.5, -1.5):
        lens.add_field(y=fy)
    lens.add_wavelength(value=0.486)
    lens.add_wavelength(value=0.587, is_primary=True)
    lens.add_wavelength(value=0.656)
    lens.add_surface(index=0, radius=INF, thickness=INF)
    lens.add_surface(index=1, radius=-99.9773, thickness=-20.0003, conic=2.35645e-06, material="mirror",
                     rx=-0.261801, is_stop=True, surface_type="polynomial", coefficients=to_array(TMA_COEFFS[0]))
    lens.add_surface(index=2, radius=-100.019, thickness=20.0003, conic=-2.21455e-07, material="mirror",
                     rx=-0.174532, dy=-11.5, surface_type="polynomial", coefficients=to_array(TMA_COEFFS[1]))
    lens.add_surface(index=3, radius=-99.9774, thickness=-20.0003, conic=5.32063e-08, material="mirror",
                     rx=-0.0174547, dy=-15, surface_type="polynomial", coefficients=to_array(TMA_COEFFS[2]))
    lens.add_surface(index=4, dy=-19.3)
    return lens


ZERNIKE_PLATE_COEFFS = [0.0, 0.0, 0.0, 2.0e-2, -1.5e-2, 1.0e-2, 4.0e-3, -3.0e-3, 2.5e-3, 0.0, 1.0e-3, -8.0e-4]


def build_freeform_plate(lens, ap, zernike_type="fringe"):
    """Refractive test system exercising what the sample lenses do not: a Zernike freeform (fringe / noll /
    standard), an odd asphere, tilts about all three axes with decenters, elliptical / rectangular / boolean
    apertures, an absorbing glass, and a steep glass-to-air exit that totally reflects part of the pupil."""
    lens.set_aperture(aperture_type="EPD", value=12)
    lens.set_field_type(field_type="angle")
    for fy in (0, 6):
        lens.add_field(y=fy)
    lens.add_wavelength(value=0.55, is_primary=True)
    lens.add_surface(index=0, radius=INF, thickness=INF)
    lens.add_surface(index=1, radius=60.0, thickness=6.0, material="N-SK16", is_stop=True, surface_type="zernike",
                     conic=-0.3, coefficients=list(ZERNIKE_PLATE_COEFFS), zernike_type=zernike_type, norm_radius=9.5,
                     aperture=ap.RadialAperture(r_max=7.0, r_min=0.5))
    lens.add_surface(index=2, radius=-45.0, thickness=4.0, surface_type="odd_asphere", conic=0.1,
                     coefficients=[0.0, 1.0e-4, -2.0e-5, 3.0e-6], rx=0.03, ry=-0.02, dx=0.4,
                     aperture=ap.DifferenceAperture(ap.EllipticalAperture(6.5, 5.5, 0.2, -0.1),
                                                    ap.RectangularAperture(-1.0, 1.0, -0.6, 0.6)))
    lens.add_surface(index=3, radius=35.0, thickness=5.0, material=("F5", "schott"), conic=0.0, rz=0.5, ry=0.04,
                     dy=-0.3, aperture=ap.UnionAperture(ap.OffsetRadialAperture(5.0, 0.0, 0.5, 0.5),
                                                        ap.IntersectionAperture(ap.RadialAperture(6.0),
                                                                                ap.RectangularAperture(-6, 6, -2, 2))))
    lens.add_surface(index=4, radius=7.5, thickness=12.0, conic=-0.2)        # steep exit: TIR at the rim
    lens.add_surface(index=5)
    return lens


def build_anamorphic(lens):
    """Anamorphic test system for the SURVEY.md 8f N4 geometries: two biconics (one with a flat x profile), a
    tilted toroid with a conic + polynomial Y-Z profile and a cylinder (toroid with an infinite rotation radius)."""
    lens.set_aperture(aperture_type="EPD", value=14)
    lens.set_field_type(field_type="angle")
    for fx, fy in ((0, 0), (0, 5), (3, 4)):
        lens.add_field(y=fy, x=fx)
    lens.add_wavelength(value=0.55, is_primary=True)
    lens.add_surface(index=0, radius=INF, thickness=INF)
    lens.add_surface(index=1, surface_type="biconic", radius_x=40.0, radius_y=55.0, conic_x=-0.5, conic_y=0.3,
                     thickness=5.0, material="N-BK7", is_stop=True)
    lens.add_surface(index=2, surface_type="biconic", radius_x=INF, radius_y=-80.0, conic_y=-1.2, thickness=3.0)
    lens.add_surface(index=3, surface_type="toroidal", radius_x=30.0, radius_y=60.0, conic=-0.4,
                     toroidal_coeffs_poly_y=[1.0e-5, -2.0e-8], thickness=4.0, material="N-SK16", rx=0.02, dy=0.2)
    lens.add_surface(index=4, surface_type="toroidal", radius_x=INF, radius_y=-50.0, thickness=30.0)
    lens.add_surface(index=5)
    return lens


CHEBYSHEV_COEFFS = [[0.0, 2.0e-2, -1.0e-2], [1.5e-2, 0.0, 4.0e-3], [-6.0e-3, 3.0e-3, 1.0e-3]]


def build_chebyshev_plate(lens, to_array=list):
    """A Chebyshev freeform front surface (3 x 3 coefficients, two of them zero) on a singlet."""
    lens.set_aperture(aperture_type="EPD", value=10)
    lens.set_field_type(field_type="angle")
    for fy in (0, 5):
        lens.add_field(y=fy)
    lens.add_wavelength(value=0.55, is_primary=True)
    lens.add_surface(index=0, radius=INF, thickness=INF)
    lens.add_surface(index=1, surface_type="chebyshev", radius=80.0, conic=-0.2, thickness=5.0, material="N-BK7",
                     is_stop=True, coefficients=to_array(CHEBYSHEV_COEFFS), norm_x=12.0, norm_y=12.0)
    lens.add_surface(index=2, radius=-70.0, thickness=40.0)
    lens.add_surface(index=3)
    return lens


def build_coated_relay(lens, coating_cls):
    """Paraxial thin lenses (one coated), SimpleCoating on refracting surfaces and on a mirror: the interaction
    models / coatings of SURVEY.md 8f N4. The surface after each thin lens renormalises the direction.
    The coatings are attached after construction: the reference's add_surface replaces any coating passed to it by
    a FresnelCoating when the surface is linked to its predecessor (surfaces/standard_surface.py:100-108)."""
    lens.set_aperture(aperture_type="EPD", value=8)
    lens.set_field_type(field_type="angle")
    for fy in (0, 4):
        lens.add_field(y=fy)
    lens.add_wavelength(value=0.55, is_primary=True)
    lens.add_surface(index=0, radius=INF, thickness=INF)
    lens.add_surface(index=1, surface_type="paraxial", f=50.0, thickness=20.0, is_stop=True)
    lens.add_surface(index=2, radius=40.0, thickness=4.0, material="N-BK7")
    lens.add_surface(index=3, radius=-60.0, conic=-0.5, thickness=6.0)
    lens.add_surface(index=4, surface_type="paraxial", f=-120.0, thickness=8.0)
    lens.add_surface(index=5, radius=-200.0, thickness=-10.0, material="mirror")
    lens.add_surface(index=6)
    for k, (t, r) in {2: (0.96, 0.04), 3: (0.97, 0.03), 4: (0.9, 0.1), 5: (0.05, 0.92)}.items():
        lens.surface_group.surfaces[k].interaction_model.coating = coating_cls(t, r)
    return lens


class Anamorphic(Optic):
    def __init__(self):
        super().__init__("Anamorphic")
        build_anamorphic(self)
        self.set_paraxial(EPL=0.0, EPD=14.0)


class CoatedRelay(Optic):
    def __init__(self):
        super().__init__("CoatedRelay")
        from .surfaces import SimpleCoating
        build_coated_relay(self, SimpleCoating)
        self.set_paraxial(EPL=0.0, EPD=8.0)


class ChebyshevPlate(Optic):
    def __init__(self):
        super().__init__("ChebyshevPlate")
        build_chebyshev_plate(self)
        self.set_paraxial(EPL=0.0, EPD=10.0)


class ThreeMirrorAnastigmat(Optic):
    def __init__(self):
        super().__init__("TMA")
        build_three_mirror(self)
        self.set_paraxial(EPL=0.0, EPD=10.0)


class FreeformPlate(Optic):
    def __init__(self, zernike_type="fringe"):
        super().__init__("FreeformPlate")
        from . import apertures
        build_freeform_plate(self, apertures, zernike_type)
        self.set_paraxial(EPL=0.0, EPD=12.0)
