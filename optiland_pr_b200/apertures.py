"""Physical apertures: parameter holders mirroring optiland/physical_apertures/*.py.

`contains` / `clip` (base.py:71-82, radial.py:56-70, offset_radial.py:48, elliptical.py:42,
rectangular.py:42, boolean trees base.py:255-336) are compiled by prescription.py into a
reverse-Polish program that the kernel evaluates on the local x, y of every ray.
"""
from __future__ import annotations

from numbers import Number


class BaseAperture:
    def __or__(self, other):
        return UnionAperture(self, other)

    def __add__(self, other):
        return UnionAperture(self, other)

    def __and__(self, other):
        return IntersectionAperture(self, other)

    def __sub__(self, other):
        return DifferenceAperture(self, other)


class RadialAperture(BaseAperture):
    def __init__(self, r_max, r_min=0):
        self.r_max = r_max
        self.r_min = r_min


class OffsetRadialAperture(RadialAperture):
    def __init__(self, r_max, r_min=0, offset_x=0, offset_y=0):
        super().__init__(r_max, r_min)
        self.offset_x = offset_x
        self.offset_y = offset_y


class EllipticalAperture(BaseAperture):
    def __init__(self, a, b, offset_x=0, offset_y=0):
        self.a = a
        self.b = b
        self.offset_x = offset_x
        self.offset_y = offset_y


class RectangularAperture(BaseAperture):
    def __init__(self, x_min, x_max, y_min, y_max):
        self.x_min = x_min
        self.x_max = x_max
        self.y_min = y_min
        self.y_max = y_max


class BaseBooleanAperture(BaseAperture):
    def __init__(self, a, b):
        self.a = a
        self.b = b


class UnionAperture(BaseBooleanAperture):
    pass


class IntersectionAperture(BaseBooleanAperture):
    pass


class DifferenceAperture(BaseBooleanAperture):
    pass


def configure_aperture(aperture):
    """radial.py:16-29: a scalar is a clear DIAMETER."""
    if aperture is None:
        return None
    if isinstance(aperture, Number):
        return RadialAperture(r_max=aperture / 2)
    if isinstance(aperture, BaseAperture):
        return aperture
    raise ValueError(f"Invalid `aperture` provided: {aperture}. Must be scalar or of type `BaseAperture`.")
