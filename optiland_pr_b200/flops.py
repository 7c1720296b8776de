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


_GEOM_NAMES = {L.GEOM_PLANE: "plane", L.GEOM_STANDARD: "conic", L.GEOM_EVEN_ASPHERE: "even asphere",
               L.GEOM_ODD_ASPHERE: "odd asphere", L.GEOM_POLYNOMIAL: "polynomial", L.GEOM_ZERNIKE: "zernike",
               L.GEOM_BICONIC: "biconic", L.GEOM_TOROIDAL: "toroidal", L.GEOM_CHEBYSHEV: "chebyshev"}


def flops_per_ray(packed, nr_iters=None) -> float:
    """Algorithmic forward flops for one ray through every intersected surface."""
    return sum(row["flops"] for row in flop_model(packed, nr_iters))


def flop_model(packed, nr_iters=None) -> list:
    """The enumerated model, one row per intersected surface ("the builder's enumerated model must be
    printed with the result"): geometry, coefficient terms m, Newton updates K, active rotations, aperture leaves,
    absorbing medium, and the flops they add up to."""
    iters = nr_iters if nr_iters is not None else packed.nr_iters()
    rows = []
    for k, lay in enumerate(packed.layouts):
        m = 0
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
        rows.append({"surface": int(lay.index), "geometry": _GEOM_NAMES.get(geom, str(geom)), "mirror": bool(mirror),
                     "terms_m": int(m), "newton_K": int(iters[k]), "rotations": bin(fflags).count("1"),
                     "aperture_leaves": sum(1 for q in range(int(packed.ops[row + L.OP_APER_LEN]))
                                            if int(packed.ops[int(packed.ops[row + L.OP_APER_OFF]) + q]) < 10),
                     "absorbing": bool(flags & L.F_ABSORB), "flops": f})
    return rows


def bytes_per_ray(itemsize: int, n_wavelengths: int, n_record: int = 0, opd_in: bool = False,
                  dir0: bool = False) -> float:
    """Algorithmic HBM bytes for one ray, forward: 7 (+1 wavelength when W > 1, +1 initial OPD)
    arrays read, 8 written, + 8 per recorded surface, + 3 for L0 M0 N0."""
    n_in = 7 + (1 if n_wavelengths > 1 else 0) + (1 if opd_in else 0)
    return float(itemsize * (n_in + 8 + 8 * n_record + (3 if dir0 else 0)))
