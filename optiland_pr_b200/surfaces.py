"""Surface / ObjectSurface / SurfaceGroup mirroring the reference's interface for the hot path.

    SurfaceGroup.trace(rays, skip=0)   optiland/surfaces/surface_group.py:232-244
    SurfaceGroup.x/y/z/L/M/N/intensity/opd  (stack of non-empty records)  :95-140
    SurfaceGroup.add_surface           :246-353 (+ factories/*.py for cs / material / geometry)
    Surface._record / reset            optiland/surfaces/standard_surface.py:244-286

`trace` performs ONE launch of the fused CUDA kernel for the whole surface loop; the
per-surface recorded arrays the analyses read afterwards are views into the kernel's record
buffer. There is no per-surface Python path and no CPU fallback in this package: a surface the
kernel does not implement raises `NotEligible`.
"""
from __future__ import annotations

import math

import torch

from . import backend as be
from .apertures import configure_aperture
from .coordinate_system import CoordinateSystem
from .geometries import (BiconicGeometry, ChebyshevPolynomialGeometry, EvenAsphere, OddAsphere, Plane,
                         PolynomialGeometry, StandardGeometry, ToroidalGeometry, ZernikePolynomialGeometry)
from .materials import BaseMaterial, IdealMaterial, Material
from .prescription import PackedSystem, pack_surface_group
from .rays import RealRays
from .trace import TraceResult, fused_trace

_RECORD_FIELDS = ("x", "y", "z", "L", "M", "N", "intensity", "opd")


class RefractiveReflectiveModel:
    """interactions/refractive_reflective_model.py:21-55 (coatings / BSDF out of scope)."""

    interaction_type = "refractive_reflective"

    def __init__(self, parent_surface=None, is_reflective=False, coating=None, bsdf=None):
        self.parent_surface = parent_surface
        self.is_reflective = is_reflective
        self.coating = coating
        self.bsdf = bsdf


class ThinLensInteractionModel(RefractiveReflectiveModel):
    """interactions/thin_lens_interaction_model.py:23-113: a paraxial thin lens of focal length f on a plane."""

    interaction_type = "thin_lens"

    def __init__(self, parent_surface=None, focal_length=None, is_reflective=False, coating=None, bsdf=None):
        super().__init__(parent_surface, is_reflective, coating, bsdf)
        self.f = focal_length


class SimpleCoating:
    """coatings.py:156-224: intensity *= transmittance (refraction) or reflectance (mirror)."""

    def __init__(self, transmittance, reflectance=0):
        self.transmittance = transmittance
        self.reflectance = reflectance
        self.absorptance = 1 - reflectance - transmittance


class Surface:
    def __init__(self, previous_surface, material_post, geometry, is_stop=False, aperture=None,
                 surface_type=None, comment="", interaction_model=None):
        self.geometry = geometry
        self.previous_surface = previous_surface
        self.material_post = material_post
        self.is_stop = is_stop
        self.aperture = configure_aperture(aperture)
        self.surface_type = surface_type
        self.comment = comment
        self.interaction_model = interaction_model or RefractiveReflectiveModel(self)
        self.interaction_model.parent_surface = self
        self.thickness = 0.0
        self.reset()

    @property
    def material_pre(self):
        return (self.previous_surface.material_post if self.previous_surface is not None
                else self.material_post)

    def reset(self):
        """standard_surface.py:244-258"""
        e = torch.empty(0)
        self.x = self.y = self.z = self.L = self.M = self.N = e
        self.intensity = self.aoi = self.opd = e
        self.u = e

    def _set_record(self, rec):
        """rec: [8, N] tensor view of the kernel's record buffer."""
        for q, name in enumerate(_RECORD_FIELDS):
            setattr(self, name, rec[q])


class ObjectSurface(Surface):
    """surfaces/object_surface.py:22-72: only records the incoming rays."""

    def __init__(self, geometry, material_post, comment=""):
        super().__init__(None, material_post, geometry, comment=comment)

    @property
    def is_infinite(self):
        return math.isinf(float(self.geometry.cs.z))


class SurfaceGroup:
    def __init__(self, surfaces=None):
        self._surfaces = list(surfaces) if surfaces else []
        self._link()
        self.record = "all"      # "all" | "last" | list of surface indices
        # fp64 mode. True (default): tolerance mode -- the north-star bar, <= 1e-10 relative to the reference with
        # identical vignetting / TIR masks (reduced-algebra kernel, csrc/trace_fx.cuh). False: strict mode --
        # the reference's operation order without FMA contraction, bit-identical for + - * / sqrt, ~2x slower.
        self.fast_math = True
        self.blocked_checkpoints = True   # differentiable traces with record 'none' / 'last': tile-blocked checkpoints
        self._packed_cache = None
        self.forced_nr_iters = None  # {surface index: K}: skip the exit-rule check (sharded batches, parallel.py)
        self.last_trace = None

    # ---- structure -----------------------------------------------------------------------
    def _link(self):
        for k, s in enumerate(self._surfaces):
            s.previous_surface = self._surfaces[k - 1] if k > 0 else None

    @property
    def surfaces(self):
        return tuple(self._surfaces)

    @property
    def num_surfaces(self):
        return len(self._surfaces)

    def invalidate(self):
        """Call after mutating any prescription value in place."""
        self._packed_cache = None

    def add_surface(self, new_surface=None, surface_type="standard", comment="", index=None,
                    is_stop=False, material="air", **kwargs):
        """Subset of surface_group.py:246-353 + factories: sequential `thickness` positioning or
        absolute x/y/z, decenter dx/dy, tilts rx/ry/rz, radius, conic, coefficients, aperture."""
        if new_surface is not None:
            self._surfaces.insert(len(self._surfaces) if index is None else index, new_surface)
            self._link()
            self.invalidate()
            return new_surface
        if index is None:
            index = len(self._surfaces)
        if index != len(self._surfaces):
            raise IndexError("this mirror only appends surfaces in order")
        # coordinate system, factories/coordinate_system_factory.py:25-88
        if "z" in kwargs:
            if "thickness" in kwargs:
                raise ValueError('Cannot define both "thickness" and "z".')
            x, y, z = kwargs.get("x", 0), kwargs.get("y", 0), kwargs["z"]
        else:
            thickness = kwargs.get("thickness", 0)
            x, y = kwargs.get("dx", 0), kwargs.get("dy", 0)
            if index == 0:
                z = -thickness
            elif index == 1:
                z = 0
            else:
                prev = self._surfaces[index - 1]
                z = prev.geometry.cs.z + prev.thickness
        cs = CoordinateSystem(x=x, y=y, z=z, rx=kwargs.get("rx", 0), ry=kwargs.get("ry", 0),
                              rz=kwargs.get("rz", 0))
        # materials, factories/material_factory.py:22-93
        if isinstance(material, BaseMaterial):
            post = material
        elif isinstance(material, tuple):
            post = Material(material[0], material[1])
        elif material.lower() == "air":
            post = IdealMaterial(n=1.0, k=0.0)
        elif material.lower() == "mirror":
            post = self._surfaces[index - 1].material_post
        else:
            post = Material(material)
        # geometry, factories/geometry_factory.py:60-230, geometry_configs.py:30-183
        radius = kwargs.get("radius", math.inf)
        conic = kwargs.get("conic", 0.0)
        tol, max_iter = kwargs.get("tol", 1e-6), kwargs.get("max_iter", 100)
        coefficients = kwargs.get("coefficients")
        if surface_type == "standard":
            rinf = math.isinf(float(radius)) if not hasattr(radius, "requires_grad") else bool(torch.isinf(radius))
            geometry = Plane(cs) if rinf else StandardGeometry(cs, radius, conic)
        elif surface_type == "even_asphere":
            geometry = EvenAsphere(cs, radius, conic, tol, max_iter, coefficients)
        elif surface_type == "odd_asphere":
            geometry = OddAsphere(cs, radius, conic, tol, max_iter, coefficients)
        elif surface_type == "polynomial":
            geometry = PolynomialGeometry(cs, radius, conic, tol, max_iter, coefficients)
        elif surface_type == "zernike":
            geometry = ZernikePolynomialGeometry(cs, radius, conic, tol, max_iter, coefficients,
                                                 kwargs.get("zernike_type", "fringe"),
                                                 kwargs.get("norm_radius", 1.0))
        elif surface_type == "paraxial":     # geometry_factory.py:352-362: a plane; the thin lens is the interaction
            geometry = Plane(cs)
        elif surface_type == "biconic":      # geometry_factory.py:229-257
            rx, ry = kwargs.get("radius_x", math.inf), kwargs.get("radius_y", math.inf)
            kx, ky = kwargs.get("conic_x", 0.0), kwargs.get("conic_y", 0.0)
            if math.isinf(float(rx)) and math.isinf(float(ry)) and kx == 0.0 and ky == 0.0:
                geometry = Plane(cs)
            else:
                geometry = BiconicGeometry(cs, rx, ry, kx, ky, tol, max_iter)
        elif surface_type == "toroidal":     # geometry_factory.py:259-279
            geometry = ToroidalGeometry(cs, kwargs.get("radius_x", math.inf), kwargs.get("radius_y", math.inf), conic,
                                        kwargs.get("toroidal_coeffs_poly_y"), tol, max_iter)
        elif surface_type == "chebyshev":    # geometry_factory.py:183-203
            geometry = ChebyshevPolynomialGeometry(cs, radius, conic, tol, max_iter, coefficients,
                                                   kwargs.get("norm_x", 1.0), kwargs.get("norm_y", 1.0))
        else:
            raise ValueError(f"surface_type {surface_type!r} is not implemented in optiland_pr_b200")
        reflective = isinstance(material, str) and material == "mirror"
        if index == 0:
            surf = ObjectSurface(geometry, post, comment)
        else:
            surf = Surface(None, post, geometry, is_stop=is_stop, aperture=kwargs.get("aperture"),
                           surface_type=surface_type, comment=comment,
                           interaction_model=(
                               ThinLensInteractionModel(None, kwargs.get("f"), reflective, kwargs.get("coating"))
                               if surface_type == "paraxial" else
                               RefractiveReflectiveModel(None, is_reflective=reflective,
                                                         coating=kwargs.get("coating"))))
        surf.thickness = kwargs.get("thickness", 0.0)
        self._surfaces.append(surf)
        self._link()
        self.invalidate()
        return surf

    # ---- recorded arrays, surface_group.py:95-140 -------------------------------------------
    def _stacked(self, name):
        return torch.stack([getattr(s, name) for s in self._surfaces if getattr(s, name).numel() > 0])

    x = property(lambda self: self._stacked("x"))
    y = property(lambda self: self._stacked("y"))
    z = property(lambda self: self._stacked("z"))
    L = property(lambda self: self._stacked("L"))
    M = property(lambda self: self._stacked("M"))
    N = property(lambda self: self._stacked("N"))
    intensity = property(lambda self: self._stacked("intensity"))
    opd = property(lambda self: self._stacked("opd"))

    def reset(self):
        for s in self._surfaces:
            s.reset()

    def _finish(self, rays, packed, out, record, dir0, nr_iters):
        """Hand the kernel's buffers to the objects the analyses read (surface_group.py:95-140,
        standard_surface.py:266-286, real_rays.py:82-86)."""
        # ObjectSurface.trace: record only, object_surface.py:56-72. The launch state is copied (the reference
        # records copies) only when every surface is recorded; "last" / "none" / a list never read it.
        if self.record == "all":
            for gi in packed.object_records:
                self._surfaces[gi]._set_record(torch.stack(
                    [rays.x, rays.y, rays.z, rays.L, rays.M, rays.N, rays.i, rays.opd]))
        for slot, gi in enumerate(packed.record_surfaces):
            self._surfaces[gi]._set_record(record[slot] if record is not None else out)
        rays.x, rays.y, rays.z, rays.L, rays.M, rays.N, rays.i, rays.opd = (out[q] for q in range(8))
        rays.L0, rays.M0, rays.N0 = dir0[0], dir0[1], dir0[2]
        self._packed_cache = {packed.layouts[k].index: nr_iters[k] for k in packed.nr_rows()}

    def _wants_grad_params(self) -> bool:
        if not torch.is_grad_enabled():
            return False
        if be.grad_mode.requires_grad:
            return True
        for s in self._surfaces:
            g = s.geometry
            vals = [g.cs.x, g.cs.y, g.cs.z, g.cs.rx, g.cs.ry, g.cs.rz, getattr(g, "radius", None), getattr(g, "k", None)]
            co = getattr(g, "coefficients", None)
            if isinstance(co, torch.Tensor):
                vals.append(co)
            else:
                for c in (co if co is not None else []):
                    vals += list(c) if isinstance(c, (list, tuple)) else [c]
            if any(isinstance(v, torch.Tensor) and v.requires_grad for v in vals):
                return True
        return False

    def _wants_grad(self, rays) -> bool:
        """Autograd is recorded when grad mode is on (torch_backend.py:45-64) or when any ray component or
        prescription entry is a tensor that requires grad."""
        if not torch.is_grad_enabled():
            return False
        if be.grad_mode.requires_grad:
            return True
        if any(isinstance(t, torch.Tensor) and t.requires_grad for t in rays.components() + [rays.opd]):
            return True
        return self._wants_grad_params()

    # ---- the hot path ----------------------------------------------------------------------
    def pack(self, wavelengths, skip=0) -> PackedSystem:
        """Pack on every call (microseconds; prescriptions are mutable) but remember the Newton
        trip counts K_s that the last trace of this group settled on as the next guess."""
        packed = pack_surface_group(self._surfaces, wavelengths, record=self.record, skip=skip)
        guess = self.forced_nr_iters or self._packed_cache or {}
        for k in packed.nr_rows():
            lay = packed.layouts[k]
            packed.ops[k * 16 + 6] = min(lay.max_iter, guess.get(lay.index, 2))
        return packed

    def trace(self, rays, skip=0, wavelengths=None):
        """surface_group.py:232-244. `wavelengths` (distinct values of rays.w) may be passed to
        avoid the device->host sync of torch.unique."""
        if not isinstance(rays, RealRays):
            raise TypeError("optiland_pr_b200.SurfaceGroup.trace only traces RealRays")
        self.reset()
        if wavelengths is None:
            wavelengths = torch.unique(rays.w).cpu().tolist()
        ins = rays.components()
        wants_grad = self._wants_grad(rays)
        packed = None if wants_grad else self.pack(wavelengths, skip)   # the autograd path packs what it needs
        if wants_grad:
            # one autograd node: forward kernel (records = checkpoints) + backward kernel (autograd.py)
            from .autograd import prescription_tensor, trace_differentiable
            def repack(mode):
                pk = pack_surface_group(self._surfaces, wavelengths, record=mode, skip=skip)
                guess = self.forced_nr_iters or self._packed_cache or {}
                for k in pk.nr_rows():
                    pk.ops[k * 16 + 6] = min(pk.layouts[k].max_iter, guess.get(pk.layouts[k].index, 2))
                return pk
            out = None
            if self.record in ("none", "last", None) and self.fast_math and self.blocked_checkpoints:
                # nobody reads per-surface records: tile-blocked checkpoints (56 B per ray and surface, no second
                # copy of the final state) instead of [S, 8, N] records -- lenses of planes / conics / even aspheres
                pk = repack("none")
                # (a lens whose last trace needed more Newton updates than the reduced-algebra adjoint unrolls goes
                # straight to the record path instead of tracing twice)
                if pk.reduced_adjoint() and max(pk.nr_iters(), default=0) <= 3:
                    presc_t = prescription_tensor(pk, self._surfaces, ins[0].dtype, ins[0].device)
                    try:
                        out, record, dir0 = trace_differentiable(pk, presc_t, ins, opd_in=rays.opd, fast=True,
                                                                 blocked=True)
                        packed, record = pk, None
                    except NotImplementedError:     # e.g. a surface that needs more than 3 Newton updates
                        out = None
            if out is None:
                packed = repack("all")
                presc_t = prescription_tensor(packed, self._surfaces, ins[0].dtype, ins[0].device)
                out, record, dir0 = trace_differentiable(packed, presc_t, ins, opd_in=rays.opd, fast=self.fast_math)
            res = TraceResult(out=out, record=record, dir0=dir0, nr_iters=packed.nr_iters(), status=0, launches=1)
        else:
            res = fused_trace(packed, ins, opd_in=rays.opd, want_dir0=True, fast=self.fast_math,
                              verify_nr=self.forced_nr_iters is None)
        if res.status & 1:
            raise ValueError("Zernike coordinates must be normalized to [-1, 1]. Consider updating "
                             "the normalization radius to 1.1x the surface aperture.")
        if res.status & 8:
            raise ValueError("Chebyshev input coordinates must be normalized to [-1, 1]. Consider updating the "
                             "normalization factors.")
        if res.status & 2:
            raise ValueError("a ray wavelength is missing from `wavelengths`")
        self._finish(rays, packed, res.out, res.record, res.dir0, res.nr_iters)
        if self.record == "last" and not packed.record_surfaces:   # blocked checkpoints: the final state IS the record
            self._surfaces[-1]._set_record(res.out)
        self.last_trace = res
        return rays
