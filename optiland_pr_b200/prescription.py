"""Prescription packer: SurfaceGroup -> flat tables for libolrt (layout in include/olrt.h).

Walks `surface_group.surfaces` exactly as `SurfaceGroup.trace` does
(optiland/surfaces/surface_group.py:232-244) and turns every `Surface` into one row of the
int32 `ops` table plus one block of reals. It is duck-typed on the reference's attribute
names (geometry.radius / .k / .coefficients / .cs.{x,y,z,rx,ry,rz,reference_cs},
material_post.n(w) / .k(w), aperture.r_max ..., interaction_model.is_reflective), so the
same code packs the reference's own objects (drop-in adapter) and this package's mirror
classes.

Refractive index n and extinction k are evaluated here, once per (surface, distinct
wavelength), with the surface's own material objects (optiland/materials/base.py:73-119),
and staged to the kernel -- the reference re-evaluates them per ray array on every call.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np

from . import _lib as L

_GEOM_BY_NAME = {
    "Plane": L.GEOM_PLANE,
    "StandardGeometry": L.GEOM_STANDARD,
    "EvenAsphere": L.GEOM_EVEN_ASPHERE,
    "OddAsphere": L.GEOM_ODD_ASPHERE,
    "PolynomialGeometry": L.GEOM_POLYNOMIAL,
    "ZernikePolynomialGeometry": L.GEOM_ZERNIKE,
    "BiconicGeometry": L.GEOM_BICONIC,
    "ToroidalGeometry": L.GEOM_TOROIDAL,
    "ChebyshevPolynomialGeometry": L.GEOM_CHEBYSHEV,
}
_NR_GEOMS = (L.GEOM_EVEN_ASPHERE, L.GEOM_ODD_ASPHERE, L.GEOM_POLYNOMIAL, L.GEOM_ZERNIKE, L.GEOM_BICONIC,
             L.GEOM_TOROIDAL, L.GEOM_CHEBYSHEV)


def _curvature(radius):
    """be.where(be.isinf(R) | (R == 0), 0.0, 1.0 / R), biconic.py:64-65 / toroidal.py:77-79."""
    r = np.float64(radius)
    return np.float64(0.0) if (np.isinf(r) or r == 0) else np.float64(1.0) / r


class NotEligible(Exception):
    """The surface group contains something the fused kernel does not implement."""


# values of the scalar torch parameters of the lens being packed, read back in ONE device->host copy
# (prefetch_tensor_scalars): id(tensor) -> (tensor, float). Without it every tensor-valued radius / conic / thickness is
# its own synchronising read-back, and an optimisation step packs the lens several times.
_TENSOR_VALUES: dict = {}


def prefetch_tensor_scalars(surfaces) -> None:
    """Collect every scalar CUDA tensor among the packer's geometry inputs and fetch their values with one copy."""
    import torch
    found = []
    for s in surfaces:
        g = getattr(s, "geometry", None)
        if g is None:
            continue
        cs = getattr(g, "cs", None)
        vals = [getattr(g, "radius", None), getattr(g, "k", None), getattr(s, "thickness", None)]
        if cs is not None:
            vals += [getattr(cs, nm, None) for nm in ("x", "y", "z", "rx", "ry", "rz")]
        co = getattr(g, "coefficients", None)
        if isinstance(co, (list, tuple)):
            vals += list(co)
        for v in vals:
            if isinstance(v, torch.Tensor) and v.is_cuda and v.numel() == 1:
                found.append(v)
    _TENSOR_VALUES.clear()
    if not found:
        return
    flat = torch.stack([v.detach().reshape(()).to(torch.float64) for v in found]).cpu().tolist()
    for v, x in zip(found, flat):
        _TENSOR_VALUES[id(v)] = (v, x)


def torch_float64():
    import torch
    return torch.float64


def _f(v) -> float:
    """Scalar value of a python number / numpy 0-d or 1-element array / torch tensor."""
    if type(v) is float or type(v) is int:     # the common case: no NumPy round trip
        return float(v)
    if hasattr(v, "detach"):
        hit = _TENSOR_VALUES.get(id(v))
        if hit is not None and hit[0] is v:
            return hit[1]
        return float(v.detach().cpu().reshape(-1)[0].item())
    a = np.asarray(v, dtype=np.float64).reshape(-1)
    return float(a[0])


def _sq(v):
    """`v**2` evaluated on the object itself, the way the reference's `contains` does."""
    return _f(v**2)


def zernike_radial_coefficients(n: int, m: int) -> list[float]:
    """Coefficients of R_n^|m|(rho), highest power first (zernike/base.py:215-239)."""
    ma = abs(m)
    out = []
    for k in range((n - ma) // 2 + 1):
        num = math.factorial(n - k)
        den = math.factorial(k) * math.factorial((n + ma) // 2 - k) * math.factorial((n - ma) // 2 - k)
        out.append(float((-1) ** k * num // den) if num % den == 0 else (-1) ** k * num / den)
    return out


@dataclass
class SurfaceLayout:
    """Where one surface's parameters sit inside a prescription (offsets in reals)."""

    index: int            # index into surface_group.surfaces
    geom: int
    doff: int             # start of the block
    frames: int           # offset of frame 0
    nframes: int
    radius: int
    conic: int
    tol: int
    norm_radius: int
    media: int            # n_pre[W], n_post[W], k_pre[W]
    coef: int
    ncoef_reals: int
    aper: int
    max_iter: int = 0
    tol_value: float = 0.0
    zernike_terms: list = field(default_factory=list)  # (n, m, offset of c, N_j)


@dataclass
class PackedSystem:
    ops: np.ndarray                 # int32 [ops_len]
    presc: np.ndarray               # float64 [n_presc, stride]
    wavelengths: list
    layouts: list                   # SurfaceLayout per intersected surface
    record_surfaces: list           # surface_group index per record slot
    object_records: list            # surface_group indices of ObjectSurface (record = the input rays)
    n_surfaces: int = 0

    @property
    def stride(self) -> int:
        return int(self.presc.shape[1])

    def nr_rows(self):
        return [k for k, lay in enumerate(self.layouts) if lay.geom in _NR_GEOMS]

    def set_nr_iters(self, iters) -> None:
        for k, lay in enumerate(self.layouts):
            if lay.geom in _NR_GEOMS:
                self.ops[k * L.OPS_STRIDE + L.OP_NR_ITERS] = int(iters[k])

    def nr_iters(self):
        return [int(self.ops[k * L.OPS_STRIDE + L.OP_NR_ITERS]) for k in range(self.n_surfaces)]

    def nr_resid_words(self) -> int:
        return sum(self.nr_iters()[k] + 1 for k in self.nr_rows())

    def fastkinds(self) -> bool:
        """True when every surface is a plane / finite conic / even asphere on one untilted frame
        with no aperture or a single RadialAperture (include/olrt.h,
        OLRT_HINT_FASTKINDS; the kernel re-checks it)."""
        for k in range(self.n_surfaces):
            row = k * L.OPS_STRIDE
            op = self.ops[row:row + L.OPS_STRIDE]
            geom, flags = int(op[L.OP_GEOM]), int(op[L.OP_FLAGS])
            if op[L.OP_NFRAMES] != 1 or op[L.OP_FRAMEFLAGS] != 0:
                return False
            if flags & (L.F_COATING | L.F_THIN_LENS | L.F_RENORM):
                return False
            if geom == L.GEOM_PLANE:
                pass
            elif geom in (L.GEOM_STANDARD, L.GEOM_EVEN_ASPHERE) and not flags & L.F_RADIUS_INF:
                pass
            else:
                return False
            aoff = int(op[L.OP_APER_OFF])
            if aoff >= 0 and not (op[L.OP_APER_LEN] == 1 and self.ops[aoff] == L.AP_RADIAL):
                return False
        return True

    def differentiable(self, require_records: bool = True) -> bool:
        """True when the backward kernels cover every surface (csrc/trace_bwd.cuh: bwd_supported): plane / finite
        conic / even or odd asphere / x^i y^j polynomial (<= 32 coefficients) / Zernike freeform (<= 16 terms,
        |m| <= 12) on ONE frame (decentred, tilted or both), any aperture program, no coating or thin-lens flag, every
        surface recorded. The kernels re-check it (OLRT_ST_BAD_HINT)."""
        if require_records and self.record_surfaces != [lay.index for lay in self.layouts]:
            return False
        for k in range(self.n_surfaces):
            row = k * L.OPS_STRIDE
            op = self.ops[row:row + L.OPS_STRIDE]
            geom, flags = int(op[L.OP_GEOM]), int(op[L.OP_FLAGS])
            if op[L.OP_NFRAMES] != 1:
                return False
            if flags & (L.F_COATING | L.F_THIN_LENS | L.F_RENORM):
                return False
            if geom == L.GEOM_PLANE:
                continue
            if geom not in (L.GEOM_STANDARD, L.GEOM_EVEN_ASPHERE, L.GEOM_ODD_ASPHERE, L.GEOM_POLYNOMIAL,
                            L.GEOM_ZERNIKE) or flags & L.F_RADIUS_INF:
                return False
            if geom == L.GEOM_ZERNIKE and any(abs(m) > L.BWG_MAXM for (_, m, _, _) in self.layouts[k].zernike_terms):
                return False
        return self.coef_slots() <= L.BWG_MAXC

    def coef_slots(self) -> int:
        """Coefficient gradient slots per surface the general adjoint kernel needs (a Zernike term takes two: its c
        and its c N entry); travels in bits 8-15 of olrt_sys.kernel_hint."""
        need = 0
        for k in range(self.n_surfaces):
            row = k * L.OPS_STRIDE
            geom, nc = int(self.ops[row + L.OP_GEOM]), int(self.ops[row + L.OP_NCOEF])
            if geom in (L.GEOM_PLANE, L.GEOM_STANDARD):
                continue
            need = max(need, 2 * nc if geom == L.GEOM_ZERNIKE else nc)
        return need

    def reduced_adjoint(self) -> bool:
        """True when the reduced-algebra adjoint (csrc/trace_bx.cuh) covers the lens: the fast kinds with <= 8
        even-asphere coefficients per surface. Other differentiable lenses take the general kernel."""
        if not self.fastkinds():
            return False
        return all(int(self.ops[k * L.OPS_STRIDE + L.OP_NCOEF]) <= L.BX_MAXC
                   for k in range(self.n_surfaces)
                   if int(self.ops[k * L.OPS_STRIDE + L.OP_GEOM]) == L.GEOM_EVEN_ASPHERE)

    def ordinary(self) -> bool:
        """fastkinds() and every surface centred (frame x = y = +0.0), refracting, without an aperture
        (include/olrt.h, OLRT_HINT_ORDINARY). Judged on this object's own prescription rows."""
        if not self.fastkinds():
            return False
        pz = np.float64(0.0).tobytes()
        for k, lay in enumerate(self.layouts):
            row = k * L.OPS_STRIDE
            if self.ops[row + L.OP_INTERACT] != 0 or self.ops[row + L.OP_APER_OFF] >= 0:
                return False
            for r in range(self.presc.shape[0]):
                if self.presc[r, lay.frames].tobytes() != pz or self.presc[r, lay.frames + 1].tobytes() != pz:
                    return False
        return True

    def make_sys(self, n_presc: int = 1, rays_per_presc: int = 0, own_rows: bool = True) -> "L.OlrtSys":
        """own_rows: the launch traces this object's own prescription row(s) (or a device copy with the same frame
        x / y entries) -- only then may OLRT_HINT_ORDINARY be claimed."""
        s = L.OlrtSys()
        s.kernel_hint = L.HINT_FASTKINDS if self.fastkinds() else 0
        cs_ = self.coef_slots()
        if 16 < cs_ <= L.BWG_MAXC:      # more coefficient gradient slots than the general adjoint's default
            s.kernel_hint |= cs_ << 8
        if own_rows and n_presc == self.presc.shape[0] and self.ordinary():
            s.kernel_hint |= L.HINT_ORDINARY
        if any(lay.geom == L.GEOM_ZERNIKE for lay in self.layouts):
            s.kernel_hint |= L.HINT_ZERNIKE
        s.n_surfaces = self.n_surfaces
        s.n_wavelengths = len(self.wavelengths)
        s.n_presc = n_presc
        s.presc_stride = self.stride
        s.ops_len = int(self.ops.shape[0])
        s.n_record = len(self.record_surfaces)
        s.nr_resid_words = self.nr_resid_words()
        s.rays_per_presc = rays_per_presc
        for q, w in enumerate(self.wavelengths):
            s.wavelengths[q] = float(w)
        return s


# --------------------------------------------------------------------------------------------
# materials: n(lambda), k(lambda) per distinct wavelength     optiland/materials/base.py:73-119
# --------------------------------------------------------------------------------------------
def _material_value(mat, which: str, wl: float) -> float:
    fn = getattr(mat, which)
    try:
        return _f(fn(wl))
    except Exception:
        return _f(fn(np.array([wl], dtype=np.float64)))


# --------------------------------------------------------------------------------------------
# apertures -> reverse-Polish program                         optiland/physical_apertures/*.py
# --------------------------------------------------------------------------------------------
def _compile_aperture(ap, tokens: list, params: list) -> None:
    name = type(ap).__name__
    if name == "RadialAperture":
        tokens.append(L.AP_RADIAL)
        params += [_sq(ap.r_max), _sq(ap.r_min)]
    elif name == "OffsetRadialAperture":
        tokens.append(L.AP_OFFSET_RADIAL)
        params += [_sq(ap.r_max), _sq(ap.r_min), _f(ap.offset_x), _f(ap.offset_y)]
    elif name == "EllipticalAperture":
        tokens.append(L.AP_ELLIPTICAL)
        params += [_sq(ap.a), _sq(ap.b), _f(ap.offset_x), _f(ap.offset_y)]
    elif name == "RectangularAperture":
        tokens.append(L.AP_RECTANGULAR)
        params += [_f(ap.x_min), _f(ap.x_max), _f(ap.y_min), _f(ap.y_max)]
    elif name in ("UnionAperture", "IntersectionAperture", "DifferenceAperture"):
        _compile_aperture(ap.a, tokens, params)
        _compile_aperture(ap.b, tokens, params)
        tokens.append({"UnionAperture": L.AP_OR, "IntersectionAperture": L.AP_AND,
                       "DifferenceAperture": L.AP_DIFF}[name])
    else:
        raise NotEligible(f"aperture type {name} is not implemented in the fused kernel")
    if len(tokens) > 30:
        raise NotEligible("aperture tree too deep for the fused kernel")


def _frames_of(cs) -> list:
    """reference_cs chain, root first (coordinate_system.py:73-89)."""
    chain = []
    while cs is not None:
        chain.append(cs)
        cs = getattr(cs, "reference_cs", None)
    chain.reverse()
    return chain


def _is_inf(v) -> bool:
    return math.isinf(_f(v))


def check_surface(surf) -> None:
    """Raise NotEligible unless `Surface.trace` on this surface is covered by the kernel."""
    g = surf.geometry
    gname = type(g).__name__
    if gname not in _GEOM_BY_NAME:
        raise NotEligible(f"geometry {gname} is not implemented in the fused kernel")
    im = surf.interaction_model
    if type(im).__name__ not in ("RefractiveReflectiveModel", "ThinLensInteractionModel"):
        raise NotEligible(f"interaction model {type(im).__name__} is not implemented")
    if type(im).__name__ == "ThinLensInteractionModel" and gname != "Plane":
        raise NotEligible("a thin lens on a curved geometry is not implemented")
    coating = getattr(im, "coating", None)
    if coating is not None and type(coating).__name__ != "SimpleCoating":
        raise NotEligible(f"coating {type(coating).__name__} needs polarised rays: not implemented")
    if getattr(im, "bsdf", None) is not None:
        raise NotEligible("BSDFs are not implemented in the fused kernel")
    for mat in (surf.material_pre, surf.material_post):
        pm = getattr(mat, "propagation_model", None)
        if pm is not None and type(pm).__name__ != "HomogeneousPropagation":
            raise NotEligible(f"propagation model {type(pm).__name__} is not implemented")


def media_row(packed: PackedSystem, surfaces, wavelength: float, skip: int = 0):
    """The prescription row of the same surfaces at ANOTHER wavelength, for a single-wavelength pack: only the media
    block (n_pre, n_post, k_pre per surface, optiland/materials/base.py:73-119) depends on it, so the row is a copy
    with those three entries re-evaluated -- a fraction of a full pack. Returns None when the ops table would differ
    (a medium that absorbs at one wavelength and not at the other): pack again then."""
    if len(packed.wavelengths) != 1:
        return None
    real = [s for s in list(surfaces)[skip:] if type(s).__name__ != "ObjectSurface"]
    if len(real) != len(packed.layouts):
        return None
    w = float(wavelength)
    row = packed.presc[0].copy()
    for k, (lay, surf) in enumerate(zip(packed.layouts, real)):
        n_pre = _material_value(surf.material_pre, "n", w)
        n_post = _material_value(surf.material_post, "n", w)
        k_pre = _material_value(surf.material_pre, "k", w)
        if (k_pre > 0) != bool(int(packed.ops[k * L.OPS_STRIDE + L.OP_FLAGS]) & L.F_ABSORB):
            return None
        row[lay.media], row[lay.media + 1], row[lay.media + 2] = n_pre, n_post, k_pre
    return row


def pack_surface_group(surfaces, wavelengths, record="all", skip: int = 0) -> PackedSystem:
    """Pack `surfaces[skip:]` for the distinct `wavelengths` (micrometres)."""
    wavelengths = [float(w) for w in wavelengths]
    prefetch_tensor_scalars(surfaces)
    try:
        return _pack_surface_group(surfaces, wavelengths, record, skip)
    finally:
        _TENSOR_VALUES.clear()    # values are only valid for this pack: parameters change between optimiser steps


def _pack_surface_group(surfaces, wavelengths, record, skip) -> PackedSystem:
    W = len(wavelengths)
    if not 1 <= W <= L.MAX_WAVELENGTHS:
        raise NotEligible(f"{W} distinct wavelengths (limit {L.MAX_WAVELENGTHS})")
    traced = list(enumerate(surfaces))[skip:]
    real = [(gi, s) for gi, s in traced if type(s).__name__ != "ObjectSurface"]
    object_records = [gi for gi, s in traced if type(s).__name__ == "ObjectSurface"]
    S = len(real)
    if not 1 <= S <= L.MAX_SURFACES:
        raise NotEligible(f"{S} surfaces (limit {L.MAX_SURFACES})")
    if record == "all":
        rec_set = [gi for gi, _ in real]
    elif record == "last":
        rec_set = [real[-1][0]]
    elif record in (None, "none"):
        rec_set = []
    else:
        rec_set = [int(r) for r in record]

    ops = [0] * (S * L.OPS_STRIDE)
    extra_ops: list = []
    reals: list = []
    layouts = []
    record_surfaces = []
    for k, (gi, surf) in enumerate(real):
        check_surface(surf)
        g = surf.geometry
        geom = _GEOM_BY_NAME[type(g).__name__]
        row = k * L.OPS_STRIDE
        doff = len(reals)
        frames = _frames_of(g.cs)
        if len(frames) > 10:
            raise NotEligible("reference_cs chain longer than 10 frames")
        fflags = 0
        for q, cs in enumerate(frames):
            rx, ry, rz = _f(cs.rx), _f(cs.ry), _f(cs.rz)
            reals += [_f(cs.x), _f(cs.y), _f(cs.z),
                      float(np.cos(np.float64(rz))), float(np.sin(np.float64(rz))),
                      float(np.cos(np.float64(ry))), float(np.sin(np.float64(ry))),
                      float(np.cos(np.float64(rx))), float(np.sin(np.float64(rx)))]
            fl = (1 if bool(rz) else 0) | (2 if bool(ry) else 0) | (4 if bool(rx) else 0)
            fflags |= fl << (3 * q)
        goff = len(reals)
        flags = 0
        if geom == L.GEOM_PLANE:
            radius, conic = math.inf, 0.0
        else:
            radius, conic = _f(g.radius), _f(g.k)
            if _is_inf(g.radius):
                flags |= L.F_RADIUS_INF
        tol = float(getattr(g, "tol", 0.0)) if geom in _NR_GEOMS else 0.0
        max_iter = int(getattr(g, "max_iter", 0)) if geom in _NR_GEOMS else 0
        if max_iter > L.MAX_NR_ITERS:
            raise NotEligible(f"max_iter {max_iter} > {L.MAX_NR_ITERS}")
        norm_radius = _f(getattr(g, "norm_radius", 1.0)) if geom == L.GEOM_ZERNIKE else 1.0
        reals += [radius, conic, tol, norm_radius]
        moff = len(reals)
        pre, post = surf.material_pre, surf.material_post
        n_pre = [_material_value(pre, "n", w) for w in wavelengths]
        n_post = [_material_value(post, "n", w) for w in wavelengths]
        k_pre = [_material_value(pre, "k", w) for w in wavelengths]
        if any(v > 0 for v in k_pre):  # be.any(k > 0), propagation/homogeneous.py:45-47
            flags |= L.F_ABSORB
        reals += n_pre + n_post + k_pre
        coff = len(reals)
        ncoef, aux = 0, 0
        zterms = []
        if geom in (L.GEOM_EVEN_ASPHERE, L.GEOM_ODD_ASPHERE):
            cs_ = [_f(c) for c in g.coefficients]
            ncoef = len(cs_)
            reals += cs_
        elif geom == L.GEOM_POLYNOMIAL:
            co2 = g.coefficients
            if hasattr(co2, "detach"):          # a 2-D tensor (a live optimisation variable): one copy, not one per entry
                rows = co2.detach().to("cpu", torch_float64()).tolist()
            else:
                rows = [[_f(c) for c in r] for r in co2]
            J = len(rows[0]) if rows else 0
            if any(len(r) != J for r in rows):
                raise NotEligible("ragged polynomial coefficient table")
            ncoef, aux = len(rows) * J, J
            for r in rows:
                reals += r
        elif geom == L.GEOM_BICONIC:      # cx cy (1+kx)cx^2 (1+ky)cy^2 -- the reference's own scalar expressions
            cx, cy = _curvature(_f(g.Rx)), _curvature(_f(g.Ry))
            kx, ky = np.float64(_f(g.kx)), np.float64(_f(g.ky))
            ncoef = 4
            reals += [float(cx), float(cy), float((1.0 + kx) * cx**2), float((1.0 + ky) * cy**2)]
        elif geom == L.GEOM_TOROIDAL:     # R_rot c_yz (1+k)c_yz^2 poly...
            c = _curvature(_f(g.R_yz))
            kk = np.float64(_f(g.k_yz))
            poly = [_f(v) for v in g.coeffs_poly_y]
            ncoef, aux = 3 + len(poly), len(poly)
            reals += [_f(g.R_rot), float(c), float((1.0 + kk) * c**2)] + poly
        elif geom == L.GEOM_CHEBYSHEV:    # norm_x norm_y C[i][j]
            rows = [[_f(c) for c in r] for r in np.atleast_2d(np.asarray(g.coefficients, dtype=object)).tolist()]
            J = len(rows[0]) if rows else 0
            ncoef, aux = 2 + len(rows) * J, J
            reals += [_f(g.norm_x), _f(g.norm_y)]
            for r in rows:
                reals += r
        elif geom == L.GEOM_ZERNIKE:
            z = g.zernike
            aux = S * L.OPS_STRIDE + len(extra_ops)
            coeffs = [_f(c) for c in z.coeffs]
            ncoef = len(coeffs)
            for (n, m), c in zip(z.indices, coeffs):
                n, m = int(n), int(m)
                norm = _f(z._norm_constant(n, m))
                extra_ops += [n, m]
                zterms.append((n, m, len(reals), norm))
                reals += [c, c * norm] + zernike_radial_coefficients(n, m)
        # interpreter-only extras, [focal length][coating factor] just before the aperture parameters (olrt.h)
        im = surf.interaction_model
        if type(im).__name__ == "ThinLensInteractionModel":
            flags |= L.F_THIN_LENS
            reals.append(_f(im.f))
        coating = getattr(im, "coating", None)
        if coating is not None:
            flags |= L.F_COATING
            reals.append(_f(coating.reflectance if im.is_reflective else coating.transmittance))
        if k > 0 and type(real[k - 1][1].interaction_model).__name__ == "ThinLensInteractionModel":
            flags |= L.F_RENORM
        aoff_reals = len(reals) - doff
        ap_off, ap_len = -1, 0
        if surf.aperture:
            toks: list = []
            prm: list = []
            _compile_aperture(surf.aperture, toks, prm)
            ap_off = S * L.OPS_STRIDE + len(extra_ops)
            ap_len = len(toks)
            extra_ops += toks
            reals += prm
        slot = -1
        if gi in rec_set:
            slot = len(record_surfaces)
            record_surfaces.append(gi)
        ops[row + L.OP_GEOM] = geom
        ops[row + L.OP_INTERACT] = 1 if surf.interaction_model.is_reflective else 0
        ops[row + L.OP_FLAGS] = flags
        ops[row + L.OP_DOFF] = doff
        ops[row + L.OP_NCOEF] = ncoef
        ops[row + L.OP_AUX] = aux
        ops[row + L.OP_NR_ITERS] = 0
        ops[row + L.OP_APER_OFF] = ap_off
        ops[row + L.OP_APER_LEN] = ap_len
        ops[row + L.OP_RECORD] = slot
        ops[row + L.OP_NFRAMES] = len(frames)
        ops[row + L.OP_FRAMEFLAGS] = fflags
        ops[row + L.OP_APER_DOFF] = aoff_reals
        layouts.append(SurfaceLayout(
            index=gi, geom=geom, doff=doff, frames=doff, nframes=len(frames), radius=goff,
            conic=goff + 1, tol=goff + 2, norm_radius=goff + 3, media=moff, coef=coff,
            ncoef_reals=len(reals) - coff - (len(reals) - doff - aoff_reals), aper=doff + aoff_reals,
            max_iter=max_iter, tol_value=tol, zernike_terms=zterms))
    packed = PackedSystem(
        ops=np.asarray(ops + extra_ops, dtype=np.int32),
        presc=np.asarray(reals, dtype=np.float64).reshape(1, -1),
        wavelengths=wavelengths, layouts=layouts, record_surfaces=record_surfaces,
        object_records=object_records, n_surfaces=S)
    return packed
