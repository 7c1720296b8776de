"""Surface geometries: parameter holders mirroring optiland/geometries/*.py (class names are the
reference's, so prescription.py packs either family).

    Plane                      geometries/plane.py
    StandardGeometry           geometries/standard.py:37-60      radius, k
    NewtonRaphsonGeometry      geometries/newton_raphson.py:43-61  tol, max_iter
    EvenAsphere / OddAsphere   geometries/even_asphere.py:59-76, odd_asphere.py:56-68
    PolynomialGeometry         geometries/polynomial.py:72-88
    ZernikePolynomialGeometry  geometries/zernike.py:88-131  (+ zernike/{standard,noll,fringe}.py)

distance / sag / surface_normal are NOT here: they are device code (csrc/trace_common.cuh,
csrc/trace_geom.cuh).
"""
from __future__ import annotations

import math

from .coordinate_system import CoordinateSystem


class BaseGeometry:
    def __init__(self, coordinate_system: CoordinateSystem):
        self.cs = coordinate_system


class Plane(BaseGeometry):
    def __init__(self, coordinate_system):
        super().__init__(coordinate_system)
        self.radius = math.inf
        self.is_symmetric = True

    def flip(self):
        pass


class StandardGeometry(BaseGeometry):
    def __init__(self, coordinate_system, radius, conic=0.0):
        super().__init__(coordinate_system)
        self.radius = radius
        self.k = conic
        self.is_symmetric = True

    def flip(self):
        self.radius = -self.radius


class NewtonRaphsonGeometry(StandardGeometry):
    def __init__(self, coordinate_system, radius, conic=0.0, tol=1e-10, max_iter=100):
        super().__init__(coordinate_system, radius, conic)
        self.tol = tol
        self.max_iter = max_iter


class EvenAsphere(NewtonRaphsonGeometry):
    def __init__(self, coordinate_system, radius, conic=0.0, tol=1e-10, max_iter=100, coefficients=None):
        super().__init__(coordinate_system, radius, conic, tol, max_iter)
        self.coefficients = list(coefficients) if coefficients is not None else []
        self.order = 2


class OddAsphere(EvenAsphere):
    def __init__(self, coordinate_system, radius, conic=0.0, tol=1e-10, max_iter=100, coefficients=None):
        super().__init__(coordinate_system, radius, conic, tol, max_iter, coefficients)
        self.order = 1


class PolynomialGeometry(NewtonRaphsonGeometry):
    def __init__(self, coordinate_system, radius, conic=0.0, tol=1e-10, max_iter=100, coefficients=None):
        super().__init__(coordinate_system, radius, conic, tol, max_iter)
        rows = coefficients if coefficients is not None else []
        rows = [list(r) for r in rows] if len(rows) and hasattr(rows[0], "__len__") else ([list(rows)] if len(rows) else [])
        if not rows:
            rows = [[0.0]]
        self.coefficients = rows
        self.is_symmetric = False


class BiconicGeometry(NewtonRaphsonGeometry):
    """geometries/biconic.py:29-67: independent conic profiles in x and y; the base conic of the first Newton
    guess is the x profile (radius_x, conic_x)."""

    def __init__(self, coordinate_system, radius_x, radius_y, conic_x=0.0, conic_y=0.0, tol=1e-10, max_iter=100):
        super().__init__(coordinate_system, radius_x, conic_x, tol, max_iter)
        self.Rx, self.Ry, self.kx, self.ky = radius_x, radius_y, conic_x, conic_y
        self.is_symmetric = False

    def flip(self):
        self.Rx, self.Ry = -self.Rx, -self.Ry
        self.radius = -self.radius


class ToroidalGeometry(NewtonRaphsonGeometry):
    """geometries/toroidal.py:26-82: a Y-Z conic + even polynomial profile rotated about an axis at distance
    radius_x; the base conic of the first Newton guess is a sphere of radius_y."""

    def __init__(self, coordinate_system, radius_x, radius_y, conic=0.0, coeffs_poly_y=None, tol=1e-10,
                 max_iter=100):
        super().__init__(coordinate_system, radius_y, 0.0, tol, max_iter)
        self.R_rot, self.R_yz, self.k_yz = radius_x, radius_y, conic
        self.coeffs_poly_y = list(coeffs_poly_y) if coeffs_poly_y is not None else []
        self.is_symmetric = False

    def flip(self):
        self.R_rot, self.R_yz = -self.R_rot, -self.R_yz
        self.radius = -self.radius


class ChebyshevPolynomialGeometry(NewtonRaphsonGeometry):
    """geometries/chebyshev.py:33-98: conic + sum C[i][j] T_i(x / norm_x) T_j(y / norm_y)."""

    def __init__(self, coordinate_system, radius, conic=0.0, tol=1e-10, max_iter=100, coefficients=None,
                 norm_x=1, norm_y=1):
        super().__init__(coordinate_system, radius, conic, tol, max_iter)
        rows = coefficients if coefficients is not None else []
        rows = [list(r) for r in rows] if len(rows) and hasattr(rows[0], "__len__") else ([list(rows)] if len(rows) else [])
        self.coefficients = rows if rows else [[0.0]]
        self.norm_x, self.norm_y = norm_x, norm_y
        self.is_symmetric = False


# ---- Zernike index schemes (OSA/ANSI "standard", Noll, Fringe) --------------------------------
def _number_standard(n, m):
    return (n * (n + 2) + m) // 2


def _number_noll(n, m):
    mod = n % 4
    if (m > 0 and mod <= 1) or (m < 0 and mod >= 2):
        c = 0
    else:
        c = 1
    return n * (n + 1) // 2 + abs(m) + c


def _number_fringe(n, m):
    sgn = (m > 0) - (m < 0)
    return int((1 + (n + abs(m)) / 2) ** 2 - 2 * abs(m) + (1 - sgn) / 2)


class _ZernikeScheme:
    """coeffs + (n, m) per coefficient (zernike/base.py:27-40, :138-184)."""

    number = staticmethod(_number_standard)
    first = 0
    normalised = True

    def __init__(self, coeffs):
        self.coeffs = list(coeffs)
        self.indices = self._generate_indices(len(self.coeffs))

    @classmethod
    def _generate_indices(cls, n_indices):
        found = {}
        n = 0
        while len([j for j in found if cls.first <= j < cls.first + n_indices]) < n_indices:
            for m in range(-n, n + 1, 2):
                found[cls.number(n, m)] = (n, m)
            n += 1
            if n > 200:
                raise ValueError("zernike index generation did not terminate")
        return [found[j] for j in range(cls.first, cls.first + n_indices)]

    @classmethod
    def _norm_constant(cls, n=0, m=0):
        if not cls.normalised:
            return 1.0
        return math.sqrt((2 * n + 2) / (2 if m == 0 else 1))


class ZernikeStandard(_ZernikeScheme):
    pass


class ZernikeNoll(_ZernikeScheme):
    number = staticmethod(_number_noll)
    first = 1


class ZernikeFringe(_ZernikeScheme):
    number = staticmethod(_number_fringe)
    first = 1
    normalised = False


_ZERNIKE_TYPES = {"standard": ZernikeStandard, "noll": ZernikeNoll, "fringe": ZernikeFringe}


class ZernikePolynomialGeometry(NewtonRaphsonGeometry):
    def __init__(self, coordinate_system, radius, conic=0.0, tol=1e-10, max_iter=100, coefficients=None,
                 zernike_type="fringe", norm_radius=1.0):
        super().__init__(coordinate_system, radius, conic, tol, max_iter)
        if zernike_type not in _ZERNIKE_TYPES:
            raise ValueError("Zernike type must be one of 'standard', 'noll', or 'fringe', got "
                             f"{zernike_type}")
        if norm_radius <= 0:
            raise ValueError(f"Normalization radius must be positive, got {norm_radius}")
        self.zernike_type = zernike_type
        self.zernike = _ZERNIKE_TYPES[zernike_type](coefficients if coefficients is not None else [])
        self.norm_radius = norm_radius
        self.is_symmetric = False

    @property
    def coefficients(self):
        return self.zernike.coeffs

    @coefficients.setter
    def coefficients(self, value):
        self.zernike = _ZERNIKE_TYPES[self.zernike_type](value)
