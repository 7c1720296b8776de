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
