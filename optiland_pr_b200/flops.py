"""Algorithmic flop / byte model of the fused trace (DESIGN.md "Roofline").

Counting convention: + - * / sqrt |.| sign compare/select = 1 flop each, counted on the reference's
lines as written. Per ray-surface:
    plane refract 52, sphere/conic refract 131, sphere/conic mirror 115,
    even asphere (m coefficients, K Newton updates) 157 + 15 m + K (58 + 15 m),
    +12 per active rotation on the way in and +12 on the way out, +8 radial aperture (+8 per
    other leaf), +8 absorbing medium.
Polynomial (I x J) and Zernike (T terms) surfaces are counted with the even-asphere skeleton and
m replaced by the number of coefficient terms evaluated (I*J resp. sum over terms of 6 + radial
terms) -- stated with the result, as the survey asks.
"""
from __future__ import annotations

from . import _lib as L


def flops_per_ray(packed, nr_iters=None) -> float:
    """Algorithmic forward flops for one ray through every intersected surface."""
    iters = nr_iters if nr_iters is not None else packed.nr_iters()
    total = 0.0
    for k, lay in enumerate(packed.layouts):
        row = k * L.OPS_STRIDE
        geom = int(packed.ops[row + L.OP_GEOM])
        mirror = int(packed.ops[row + L.OP_INTERACT]) == 1
        flags = int(packed.ops[row + L.OP_FLAGS])
        if geom == L.GEOM_PLANE or (geom == L.GEOM_STANDARD and flags & L.F_RADIUS_INF):
            f = 52.0 - (16.0 if mirror else 0.0)
        elif geom == L.GEOM_STANDARD:
            f = 115.0 if mirror else 131.0
        else:
            if geom in (L.GEOM_EVEN_ASPHERE, L.GEOM_ODD_ASPHERE, L.GEOM_POLYNOMIAL):
                m = int(packed.ops[row + L.OP_NCOEF])
            else:
                m = sum(6 + (n - abs(mm)) // 2 + 1 for (n, mm, _, _) in lay.zernike_terms)
            f = 157.0 + 15.0 * m + int(iters[k]) * (58.0 + 15.0 * m) - (16.0 if mirror else 0.0)
        fflags = int(packed.ops[row + L.OP_FRAMEFLAGS])
        f += 24.0 * bin(fflags).count("1")
        f += 8.0 * sum(1 for q in range(int(packed.ops[row + L.OP_APER_LEN]))
                       if int(packed.ops[int(packed.ops[row + L.OP_APER_OFF]) + q]) < 10)
        if flags & L.F_ABSORB:
            f += 8.0
        total += f
    return total


def bytes_per_ray(itemsize: int, n_wavelengths: int, n_record: int = 0, opd_in: bool = False,
                  dir0: bool = False) -> float:
    """Algorithmic HBM bytes for one ray, forward: 7 (+1 wavelength when W > 1, +1 initial OPD)
    arrays read, 8 written, + 8 per recorded surface, + 3 for L0 M0 N0."""
    n_in = 7 + (1 if n_wavelengths > 1 else 0) + (1 if opd_in else 0)
    return float(itemsize * (n_in + 8 + 8 * n_record + (3 if dir0 else 0)))
