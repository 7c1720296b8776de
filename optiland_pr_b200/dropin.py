"""Drop-in for the UNMODIFIED reference: `install()` puts the fused sm_100a trace under the reference's own
`SurfaceGroup.trace(self, rays, skip=0)` (optiland/surfaces/surface_group.py:232-244) -- and under nothing else.

After `install()`, with `be.set_backend("torch"); be.set_device("cuda")`, everything that reaches that method --
`RealRayTracer.trace / trace_generic` (optiland/raytrace/real_ray_tracer.py:82, :122), hence `Optic.trace`,
`SpotDiagram` (analysis/spot_diagram.py:397-438), `RayOperand.*` (optimization/operand/ray.py:299-340), `Wavefront`
(wavefront/strategy.py:176-229) and the torch optimisers (optimization/optimizer/torch/base.py:95-154) -- runs on
ONE launch of `olrt_trace_fwd_*` (and `olrt_trace_bwd_*` under autograd) instead of ~150 elementwise torch kernels
per surface. The reference's objects are packed as they are (`prescription.pack_surface_group` is duck-typed on the
reference's attribute names); its paraxial code, ray generator, analyses and optimisers are untouched and stay in the
autograd graph, so d/d(parameter) also flows through the launch plane (EPL / EPD) exactly as in the reference.

Per call the stock Python loop runs instead -- never a numerical substitute -- when
  * the rays are not exactly `RealRays` (`ParaxialRays` from paraxial_ray_tracer.py:56, `PolarizedRays`),
  * the backend is not torch or the rays are not on a CUDA device (the NumPy backend is the oracle),
  * a surface is outside the kernels (`NotEligible`: BSDF, Fresnel coating, GRIN, polygon aperture, Forbes ...),
  * autograd is wanted and the lens is outside the adjoint kernels' differentiable set, or a quantity the adjoint does
    not differentiate (tilt, refractive index, polynomial / Zernike coefficient, aperture size) depends on an
    optimisation variable: the stock loop then records the graph op by op, as the reference does.
`stats()` counts which path every call took and why.
"""
from __future__ import annotations

import threading

import torch

from .prescription import NotEligible, pack_surface_group

_FIELDS = ("x", "y", "z", "L", "M", "N", "intensity", "opd")
_ZERNIKE_MSG = ("Zernike coordinates must be normalized to [-1, 1]. Consider updating the normalization radius to "
                "1.1x the surface aperture.")    # optiland/geometries/zernike.py:241-246
_CHEBYSHEV_MSG = ("Chebyshev input coordinates must be normalized to [-1, 1]. Consider updating the normalization "
                  "factors.")                    # optiland/geometries/chebyshev.py:204-217

_lock = threading.Lock()
_state = {"stock": None, "cls": None, "opts": None}
_stats: dict = {}
_NR_GUESS: dict = {}    # id(surface_group) -> {surface index: K_s the last trace settled on}


def stats(reset: bool = False) -> dict:
    """{"fused": n, "fused_grad": n, "stock:<reason>": n, ...} since install() / the last reset."""
    out = dict(_stats)
    if reset:
        _stats.clear()
    return out


def _count(key: str) -> None:
    _stats[key] = _stats.get(key, 0) + 1


def installed() -> bool:
    return _state["stock"] is not None


def _depends_on_variable(v) -> bool:
    """True for a tensor an optimiser can reach: an nn.Parameter or anything computed from one. (Under
    `be.grad_mode.enable()` every `be.array` is a requires_grad LEAF, torch_backend.py:133-147 -- those carry no
    variable and nobody reads their .grad.)"""
    return isinstance(v, torch.Tensor) and v.requires_grad and (
        v.grad_fn is not None or isinstance(v, torch.nn.Parameter))


def _undifferentiated_variable(surfaces):
    """Name of a quantity that depends on an optimisation variable but that the adjoint kernels do not differentiate
    (DESIGN.md 4.4), or None."""
    for k, s in enumerate(surfaces):
        g = getattr(s, "geometry", None)
        if g is None:
            continue
        cs = getattr(g.cs, "reference_cs", None)     # tilts of the surface's own frame are differentiated (trace_bwd.cuh)
        while cs is not None:
            for nm in ("rx", "ry", "rz", "x", "y", "z"):
                if _depends_on_variable(getattr(cs, nm, None)):
                    return f"surface {k}: {nm} of a parent frame"
            cs = getattr(cs, "reference_cs", None)
        for nm in ("norm_radius", "norm_x", "norm_y", "Rx", "Ry", "kx", "ky", "R_rot", "R_yz", "k_yz"):
            if _depends_on_variable(getattr(g, nm, None)):
                return f"surface {k}: {nm}"
        gname = type(g).__name__
        if gname not in ("EvenAsphere", "OddAsphere", "Plane", "StandardGeometry", "PolynomialGeometry",
                         "ZernikePolynomialGeometry"):
            co = getattr(g, "coefficients", None)
            if co is None and getattr(g, "zernike", None) is not None:
                co = g.zernike.coeffs
            items = [co] if isinstance(co, torch.Tensor) else list(co if co is not None else [])
            flat = []
            for c in items:
                flat += list(c) if isinstance(c, (list, tuple)) else [c]
            if any(_depends_on_variable(c) for c in flat):
                return f"surface {k}: {gname} coefficients"
        for mat in (getattr(s, "material_pre", None), getattr(s, "material_post", None)):
            for nm in ("index", "absorp"):
                if _depends_on_variable(getattr(mat, nm, None)):
                    return f"surface {k}: refractive index"
        ap = getattr(s, "aperture", None)
        if ap:
            for nm in ("r_max", "r_min", "offset_x", "offset_y", "a", "b", "x_min", "x_max", "y_min", "y_max"):
                if _depends_on_variable(getattr(ap, nm, None)):
                    return f"surface {k}: aperture {nm}"
    return None


def _wants_grad(be, rays, surfaces) -> bool:
    """Autograd is recorded when grad mode is on (torch_backend.py:45-64) or any ray / parameter tensor asks."""
    if not torch.is_grad_enabled():
        return False
    if be.grad_mode.requires_grad:
        return True
    for t in (rays.x, rays.y, rays.z, rays.L, rays.M, rays.N, rays.i, rays.opd):
        if isinstance(t, torch.Tensor) and t.requires_grad:
            return True
    for s in surfaces:
        g = getattr(s, "geometry", None)
        if g is None:
            continue
        vals = [g.cs.x, g.cs.y, g.cs.z, g.cs.rx, g.cs.ry, g.cs.rz, getattr(g, "radius", None), getattr(g, "k", None)]
        co = getattr(g, "coefficients", None)
        if isinstance(co, torch.Tensor):
            vals.append(co)
        elif isinstance(co, (list, tuple)):
            for c in co:
                vals += list(c) if isinstance(c, (list, tuple)) else [c]
        if any(isinstance(v, torch.Tensor) and v.requires_grad for v in vals):
            return True
    # a variable on a quantity the kernels do not differentiate (norm radius, index, aperture size, parent frame ...)
    # must still reach the reference's own graph
    return _undifferentiated_variable(surfaces) is not None


def _make_fused(stock, RealRays, be, opts):
    from .autograd import prescription_tensor, trace_differentiable
    from .trace import fused_trace

    def trace(self, rays, skip=0):
        """SurfaceGroup.trace (optiland/surfaces/surface_group.py:232-244) on the fused kernels."""
        if type(rays) is not RealRays:
            _count("stock:not RealRays")
            return stock(self, rays, skip)
        if be.get_backend() != "torch" or not (isinstance(rays.x, torch.Tensor) and rays.x.is_cuda):
            _count("stock:not torch/cuda")
            return stock(self, rays, skip)
        if rays.x.dtype not in (torch.float64, torch.float32) or rays.x.numel() == 0:
            _count("stock:dtype/empty")
            return stock(self, rays, skip)
        surfaces = self.surfaces
        grad = _wants_grad(be, rays, surfaces)
        if grad:
            why = _undifferentiated_variable(surfaces)
            if why is not None:
                _count("stock:grad of " + why.split(": ", 1)[1])
                return stock(self, rays, skip)
        w = rays.w
        wl = torch.unique(w.detach()).cpu().tolist() if w.numel() > 1 else [float(w.detach().reshape(-1)[0])]
        try:
            packed = pack_surface_group(surfaces, wl, record=opts["record"], skip=skip)
        except NotEligible as e:
            _count("stock:NotEligible " + str(e)[:40])
            return stock(self, rays, skip)
        if grad and not packed.differentiable(require_records=False):
            _count("stock:grad outside the adjoint kernels")
            return stock(self, rays, skip)
        guess = _NR_GUESS.get(id(self), {})
        for k in packed.nr_rows():
            lay = packed.layouts[k]
            packed.ops[k * 16 + 6] = min(lay.max_iter, guess.get(lay.index, 2))
        dtype, dev = rays.x.dtype, rays.x.device
        ins = [t.to(dtype) if t.dtype != dtype else t for t in
               (rays.x, rays.y, rays.z, rays.L, rays.M, rays.N, rays.i, rays.w)]
        n = ins[0].numel()
        ins = [t.expand(n) if t.numel() == 1 and n > 1 else t for t in ins]
        opd_in = rays.opd
        self.reset()                                                   # surface_group.py:241
        if grad:
            if packed.record_surfaces != [lay.index for lay in packed.layouts]:   # the checkpoints are the records
                packed = pack_surface_group(surfaces, wl, record="all", skip=skip)
                for k in packed.nr_rows():
                    lay = packed.layouts[k]
                    packed.ops[k * 16 + 6] = min(lay.max_iter, guess.get(lay.index, 2))
            presc = prescription_tensor(packed, surfaces, dtype, dev)
            out, record, dir0 = trace_differentiable(packed, presc, ins, opd_in=opd_in, fast=opts["fast_math"])
            status, nr = 0, packed.nr_iters()
            _count("fused_grad")
        else:
            res = fused_trace(packed, [t.detach() for t in ins], opd_in=opd_in.detach(), want_dir0=True,
                              fast=opts["fast_math"])
            out, record, dir0, status, nr = res.out, res.record, res.dir0, res.status, res.nr_iters
            _count("fused")
        if status & 1:
            raise ValueError(_ZERNIKE_MSG)
        if status & 8:
            raise ValueError(_CHEBYSHEV_MSG)
        _NR_GUESS[id(self)] = {packed.layouts[k].index: nr[k] for k in packed.nr_rows()}
        for gi in packed.object_records:                               # ObjectSurface.trace, object_surface.py:56-72
            s = surfaces[gi]
            for f, v in zip(_FIELDS, ins[:7] + [opd_in]):
                setattr(s, f, v.clone())
        for slot, gi in enumerate(packed.record_surfaces):             # Surface._record, standard_surface.py:266-286
            s = surfaces[gi]
            for q, f in enumerate(_FIELDS):
                setattr(s, f, record[slot, q])
        rays.x, rays.y, rays.z, rays.L, rays.M, rays.N, rays.i, rays.opd = (out[q] for q in range(8))
        rays.L0, rays.M0, rays.N0 = dir0[0], dir0[1], dir0[2]           # real_rays.py:82-86, :141-146
        rays.is_normalized = True
        return rays

    trace._olrt_fused = True
    return trace


def install(fast_math: bool = True, record="all"):
    """Patch the reference's `SurfaceGroup.trace`. `optiland` (the reference) must be importable.
    fast_math: tolerance mode (<= 1e-10 relative to the NumPy reference, same masks); False = bit-faithful strict mode.
    record:    "all" (the reference's behaviour: every surface keeps its post-interaction state) or "last" (only the
               image surface: `surface_group.x[-1, :]` stays valid, positive indices do not; 64 B / ray / surface
               less HBM traffic). Differentiable traces always record every surface (the records are the checkpoints).
    Returns the stock method. Loads libolrt.so first: a missing library raises here, not at the first trace."""
    from . import _lib
    _lib.load()
    import optiland.backend as be
    from optiland.rays import RealRays
    from optiland.surfaces.surface_group import SurfaceGroup
    with _lock:
        if _state["stock"] is None:
            _state["stock"] = SurfaceGroup.trace
            _state["cls"] = SurfaceGroup
        opts = {"fast_math": bool(fast_math), "record": record}
        _state["opts"] = opts
        SurfaceGroup.trace = _make_fused(_state["stock"], RealRays, be, opts)
        _stats.clear()
        _NR_GUESS.clear()
    return _state["stock"]


def uninstall() -> None:
    """Put the stock method back."""
    with _lock:
        if _state["stock"] is not None:
            _state["cls"].trace = _state["stock"]
            _state["stock"] = _state["cls"] = _state["opts"] = None
        _NR_GUESS.clear()
