"""Optic facade for the hot path: the calls analyses and operands make before reading
`surface_group.x[-1, :]`.

    Optic.add_surface / add_field / add_wavelength / set_aperture   optiland/optic/optic.py
    Optic.trace(Hx, Hy, wavelength, num_rays, distribution)         optic.py:584-609
    Optic.trace_generic(Hx, Hy, Px, Py, wavelength)                 optic.py:611-632
    RealRayTracer.trace / trace_generic                             raytrace/real_ray_tracer.py:37-133
    RayGenerator.generate_rays (non-telecentric branch)             rays/ray_generator.py:28-106
    AngleField / ObjectHeightField.get_ray_origins                  fields/field_types.py:139-169, :245-272
    FieldGroup.get_vig_factor                                       fields/field_group.py:80-109

Scope notes: ray generation covers angle fields (object at infinity or finite), object-height
fields and per-field vignetting factors; object-space telecentric systems and paraxial-image-height fields are
not mirrored. The first-order quantities EPL / EPD it needs are INPUTS (`set_paraxial`) because the paraxial
tracer (optiland/paraxial.py) is outside the hot path.
"""
from __future__ import annotations

import math

import numpy as np
import torch

from . import backend as be
from .rays import RealRays
from .surfaces import SurfaceGroup


def pupil_points(distribution: str, num_rays: int, seed=None):
    """Pupil sampling, host side (optiland/distribution.py): 'random' (:132-158, r=sqrt(U),
    theta=2*pi*U from numpy default_rng(seed)), 'hexapolar' (:189-220), 'uniform'."""
    if distribution == "random":
        rng = np.random.default_rng(seed)
        r = rng.uniform(size=num_rays)
        theta = rng.uniform(0, 2 * np.pi, size=num_rays)
        return np.sqrt(r) * np.cos(theta), np.sqrt(r) * np.sin(theta)
    if distribution == "hexapolar":
        x, y = [0.0], [0.0]
        for i in range(num_rays):
            num_theta = 6 * (i + 1)
            theta = np.linspace(0, 2 * np.pi, num_theta + 1)[:-1]
            r = (i + 1) / num_rays
            x.extend(r * np.cos(theta))
            y.extend(r * np.sin(theta))
        return np.asarray(x), np.asarray(y)
    if distribution == "uniform":
        g = np.linspace(-1, 1, num_rays)
        xx, yy = np.meshgrid(g, g)
        keep = xx**2 + yy**2 <= 1
        return xx[keep], yy[keep]
    raise ValueError(f"distribution {distribution!r} is not implemented")


def random_pupil_device(num_rays: int, seed=None, device=None, want_raw: bool = False):
    """RandomDistribution(seed).generate_points(num_rays) (optiland/distribution.py:132-158, NumPy backend generator
    = PCG64) generated ON THE DEVICE by `olrt_pupil_random_f64`: the seed's 128-bit generator state goes to the GPU,
    not 16 bytes per pupil point. The uniform draws are NumPy's bit for bit; Px, Py differ from NumPy's by the last
    bit or two of cos / sin. Returns (Px, Py) device tensors (+ raw [2, n] draws when asked)."""
    import ctypes as C

    from . import _lib as L
    lib = L.load()
    dev = torch.device(device if device is not None else be.get_device())
    # NumPy hashes the seed into the generator's 128-bit state and increment (SeedSequence): ~0.1 ms, remembered per seed
    hit = _PCG_STATE.get(seed) if isinstance(seed, int) else None
    if hit is None:
        st = np.random.default_rng(seed).bit_generator.state["state"]
        hit = (st["state"], st["inc"])
        if isinstance(seed, int):
            if len(_PCG_STATE) > 256:
                _PCG_STATE.clear()
            _PCG_STATE[seed] = hit
    mask = (1 << 64) - 1
    state = (C.c_uint64 * 2)(hit[0] >> 64, hit[0] & mask)
    inc = (C.c_uint64 * 2)(hit[1] >> 64, hit[1] & mask)
    n = int(num_rays)
    Px = torch.empty(n, dtype=torch.float64, device=dev)
    Py = torch.empty(n, dtype=torch.float64, device=dev)
    raw = torch.empty((2, n), dtype=torch.float64, device=dev) if want_raw else None
    with torch.cuda.device(dev):
        rc = lib.olrt_pupil_random_f64(C.byref(state), C.byref(inc), n, Px.data_ptr(), Py.data_ptr(),
                                       0 if raw is None else raw.data_ptr(), torch.cuda.current_stream(dev).cuda_stream)
    L.check(rc, "olrt_pupil_random_f64")
    return (Px, Py, raw) if want_raw else (Px, Py)


def hexapolar_pupil_device(num_rings: int, device=None):
    """HexagonalDistribution.generate_points(num_rings) (optiland/distribution.py:189-220) generated ON THE DEVICE by
    `olrt_pupil_hexapolar_f64`: radii and angles are NumPy's linspace values bit for bit, Px, Py within an ulp or two
    (cos / sin). Returns (Px, Py) device tensors of 1 + 3 R (R + 1) points."""
    from . import _lib as L
    lib = L.load()
    dev = torch.device(device if device is not None else be.get_device())
    R_ = int(num_rings)
    n = 1 + 3 * R_ * (R_ + 1)
    Px = torch.empty(n, dtype=torch.float64, device=dev)
    Py = torch.empty(n, dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        rc = lib.olrt_pupil_hexapolar_f64(R_, Px.data_ptr(), Py.data_ptr(), torch.cuda.current_stream(dev).cuda_stream)
    L.check(rc, "olrt_pupil_hexapolar_f64")
    return Px, Py


def uniform_pupil_device(num_points: int, device=None):
    """UniformDistribution.generate_points(num_points) (optiland/distribution.py:174-185) generated ON THE DEVICE by
    `olrt_pupil_uniform_f64`: a counting pass, a prefix sum of the row counts (the one 8-byte read-back: the number of
    kept points sizes the arrays), a writing pass. NumPy's values and mask bit for bit."""
    from . import _lib as L
    lib = L.load()
    dev = torch.device(device if device is not None else be.get_device())
    n = int(num_points)
    if n == 0:
        e = torch.empty(0, dtype=torch.float64, device=dev)
        return e, e.clone()
    counts = torch.empty(n, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        st = torch.cuda.current_stream(dev).cuda_stream
        L.check(lib.olrt_pupil_uniform_f64(n, 0, counts.data_ptr(), 0, 0, st), "olrt_pupil_uniform_f64")
        ends = torch.cumsum(counts, 0)
        total = int(ends[-1].item())
        offsets = (ends - counts).contiguous()
        Px = torch.empty(total, dtype=torch.float64, device=dev)
        Py = torch.empty(total, dtype=torch.float64, device=dev)
        if total > 0:
            L.check(lib.olrt_pupil_uniform_f64(n, offsets.data_ptr(), 0, Px.data_ptr(), Py.data_ptr(), st),
                    "olrt_pupil_uniform_f64")
    return Px, Py


_FIELD_CACHE: dict = {}
_PCG_STATE: dict = {}       # seed -> (state, inc) of numpy.random.default_rng(seed) before its first draw


class FieldTable:
    """The per-field scalars of the ray launch (include/olrt.h, olrt_pupil): `fxy` [F, 2] launch-plane offsets
    (object at infinity) or object points (finite object), `vig` [F, 2] = (vx, vy) or None, `fz` [F] or None,
    `z_start`. `to(device)` uploads them in ONE copy; `rows(f)` is the table of field f alone."""

    def __init__(self, fxy, vig, fz, z_start):
        self.fxy, self.vig, self.fz, self.z_start = fxy, vig, fz, float(z_start)
        self.fxy_d = self.vig_d = self.fz_d = None

    @property
    def n_fields(self):
        return int(self.fxy.shape[0])

    def to(self, device):
        F = self.n_fields
        parts = [self.fxy.ravel()] + ([self.vig.ravel()] if self.vig is not None else []) \
            + ([self.fz.ravel()] if self.fz is not None else [])
        host = np.ascontiguousarray(np.concatenate(parts), dtype=np.float64)
        key = (str(device), host.tobytes())
        flat = _FIELD_CACHE.get(key)     # by content: the same fields again (every analysis pass) upload nothing
        if flat is None:
            flat = torch.from_numpy(host).to(device, non_blocking=True)
            if len(_FIELD_CACHE) >= 32:
                _FIELD_CACHE.pop(next(iter(_FIELD_CACHE)))
            _FIELD_CACHE[key] = flat
        self.fxy_d = flat[:2 * F].view(F, 2)
        o = 2 * F
        if self.vig is not None:
            self.vig_d = flat[o:o + 2 * F].view(F, 2)
            o += 2 * F
        if self.fz is not None:
            self.fz_d = flat[o:o + F]
        return self

    def rows(self, f):
        t = FieldTable(self.fxy[f:f + 1], None if self.vig is None else self.vig[f:f + 1],
                       None if self.fz is None else self.fz[f:f + 1], self.z_start)
        t.fxy_d = self.fxy_d[f:f + 1]
        t.vig_d = None if self.vig_d is None else self.vig_d[f:f + 1]
        t.fz_d = None if self.fz_d is None else self.fz_d[f:f + 1]
        return t

    def kwargs(self):
        """The keyword arguments of trace.fused_trace_pupil that come from this table."""
        return dict(z_start=self.z_start, field_vig=self.vig_d, field_z=self.fz_d)


class Optic:
    def __init__(self, name=None):
        self.name = name
        self.surface_group = SurfaceGroup()
        self.fields = []          # (x, y): degrees (angle fields) or lens units (object-height fields)
        self.vignetting = []      # (vx, vy) vignetting factors per field, fields/field.py:22-32
        self.field_type = "angle"
        self.wavelengths = []     # micrometres
        self.primary_index = 0
        self.aperture = None
        self._EPL = None
        self._EPD = None
        self._XPL = None

    # ---- construction ------------------------------------------------------------------------
    def add_surface(self, **kwargs):
        return self.surface_group.add_surface(**kwargs)

    def set_aperture(self, aperture_type, value):
        self.aperture = (aperture_type, value)
        if aperture_type == "EPD":
            self._EPD = float(value)

    def set_field_type(self, field_type):
        if field_type not in ("angle", "object_height"):
            raise ValueError('field_type must be "angle" or "object_height" (paraxial image heights are not mirrored)')
        self.field_type = field_type

    def add_field(self, y=0.0, x=0.0, vx=0.0, vy=0.0):
        self.fields.append((float(x), float(y)))
        self.vignetting.append((float(vx), float(vy)))

    def get_vig_factor(self, Hx, Hy):
        """FieldGroup.get_vig_factor, fields/field_group.py:80-109: the vignetting factors of the NEAREST defined
        field (normalised coordinates). Returns (vx [n], vy [n]) NumPy arrays."""
        Hx = np.atleast_1d(np.asarray(Hx, dtype=np.float64))
        Hy = np.atleast_1d(np.asarray(Hy, dtype=np.float64))
        mf = self.max_field
        pts = np.asarray(self.fields, dtype=np.float64).reshape(-1, 2)
        if mf != 0:
            pts = pts / mf
        v = np.asarray(self.vignetting, dtype=np.float64).reshape(-1, 2)
        d2 = (Hx[:, None] - pts[None, :, 0]) ** 2 + (Hy[:, None] - pts[None, :, 1]) ** 2
        near = np.argmin(d2, axis=1)
        return v[near, 0], v[near, 1]

    def add_wavelength(self, value, is_primary=False):
        self.wavelengths.append(float(value))
        if is_primary:
            self.primary_index = len(self.wavelengths) - 1

    def set_paraxial(self, EPL, EPD=None, XPL=None):
        """Entrance-pupil location / diameter and exit-pupil location relative to the image surface
        (optiland/paraxial.py:207-280, :282-310), supplied by the caller."""
        self._EPL = float(EPL)
        if EPD is not None:
            self._EPD = float(EPD)
        if XPL is not None:
            self._XPL = float(XPL)

    # ---- OpticUpdater subset used by the optimisation variables (optic/optic_updater.py:38-142) ------
    def set_radius(self, value, surface_number):
        self.surface_group.surfaces[surface_number].geometry.radius = value
        self.surface_group.invalidate()

    def set_conic(self, value, surface_number):
        self.surface_group.surfaces[surface_number].geometry.k = value
        self.surface_group.invalidate()

    def set_asphere_coeff(self, value, surface_number, aspher_coeff_idx):
        self.surface_group.surfaces[surface_number].geometry.coefficients[aspher_coeff_idx] = value
        self.surface_group.invalidate()

    def set_thickness(self, value, surface_number):
        """optic_updater.py:67-85: every surface after `surface_number` moves by the change in thickness
        (cs.z stays a tensor expression of the thickness when `value` is a tensor)."""
        surfs = self.surface_group.surfaces
        delta = value - surfs[surface_number].thickness
        surfs[surface_number].thickness = value
        for s in surfs[surface_number + 1:]:
            s.geometry.cs.z = s.geometry.cs.z + delta
        self.surface_group.invalidate()

    @property
    def primary_wavelength(self):
        return self.wavelengths[self.primary_index]

    @property
    def max_field(self):
        """be.max(be.sqrt(x_fields**2 + y_fields**2)), fields/field_group.py:61-63."""
        if not self.fields:
            return 0.0
        f = np.asarray(self.fields, dtype=np.float64)
        return float(np.max(np.sqrt(f[:, 0] ** 2 + f[:, 1] ** 2)))

    def get_field_coords(self):
        mf = self.max_field
        return [(fx / mf if mf else 0.0, fy / mf if mf else 0.0) for fx, fy in self.fields]

    # ---- ray generation: ray_generator.py:28-106 + AngleField.get_ray_origins (infinite object) ---
    def generate_rays(self, Hx, Hy, Px, Py, wavelength) -> RealRays:
        """Hx, Hy, Px, Py: scalars, NumPy arrays or torch tensors (host tensors are copied to the
        device asynchronously -- pass pinned memory to overlap), broadcast against each other."""
        if self._EPL is None or self._EPD is None:
            raise ValueError("call set_paraxial(EPL, EPD) first")
        dt, dev = be.get_precision(), be.get_device()

        def f64(a):
            if isinstance(a, torch.Tensor):
                return a.to(device=dev, dtype=torch.float64, non_blocking=True).reshape(-1)
            return torch.as_tensor(np.asarray(a, dtype=np.float64), device=dev).reshape(-1)

        vx = vy = 1.0
        if any(v != (0.0, 0.0) for v in self.vignetting):
            hx_h = Hx.cpu().numpy() if isinstance(Hx, torch.Tensor) else np.asarray(Hx, dtype=np.float64)
            hy_h = Hy.cpu().numpy() if isinstance(Hy, torch.Tensor) else np.asarray(Hy, dtype=np.float64)
            hx_h, hy_h = np.broadcast_arrays(np.atleast_1d(hx_h), np.atleast_1d(hy_h))
            vxf, vyf = self.get_vig_factor(hx_h.reshape(-1), hy_h.reshape(-1))
            vx, vy = f64(1 - vxf), f64(1 - vyf)
        Hx, Hy, Px, Py = f64(Hx), f64(Hy), f64(Px), f64(Py)
        EPL, EPD, mf = self._EPL, self._EPD, self.max_field
        obj = self.surface_group.surfaces[0]
        if self.field_type == "object_height":
            if obj.is_infinite:
                raise ValueError('Field type cannot be "object_height" for an object at infinity.')
            x0 = (mf * Hx) * torch.ones_like(Px)
            y0 = (mf * Hy) * torch.ones_like(Px)
            z0 = torch.as_tensor(self._object_sag(x0.cpu().numpy(), y0.cpu().numpy()), device=dev) + float(obj.geometry.cs.z)
        elif obj.is_infinite:
            z_pos = [float(s.geometry.cs.z) for s in self.surface_group.surfaces[1:-1]]
            offset = EPD - min(z_pos)                              # field_types.py:223-239
            x = -torch.tan(torch.deg2rad(mf * Hx)) * (offset + EPL)
            y = -torch.tan(torch.deg2rad(mf * Hy)) * (offset + EPL)
            z = float(self.surface_group.surfaces[1].geometry.cs.z) - offset
            x0 = Px * EPD / 2 * vx + x
            y0 = Py * EPD / 2 * vy + y
            z0 = torch.full_like(x0, z)
        else:
            zo = float(obj.geometry.cs.z)
            x0 = (-torch.tan(torch.deg2rad(mf * Hx)) * (EPL - zo)) * torch.ones_like(Px)
            y0 = (-torch.tan(torch.deg2rad(mf * Hy)) * (EPL - zo)) * torch.ones_like(Px)
            z0 = torch.full_like(x0, zo)
        x1 = Px * EPD * vx / 2
        y1 = Py * EPD * vy / 2
        z1 = torch.full_like(x0, EPL)
        mag = torch.sqrt((x1 - x0) ** 2 + (y1 - y0) ** 2 + (z1 - z0) ** 2)
        is_zero = mag < 1e-9
        mag = torch.where(is_zero, torch.ones_like(mag), mag)
        L = torch.where(is_zero, torch.zeros_like(mag), (x1 - x0) / mag)
        M = torch.where(is_zero, torch.zeros_like(mag), (y1 - y0) / mag)
        N = torch.where(is_zero, torch.ones_like(mag), (z1 - z0) / mag)
        w = torch.ones_like(x0) * wavelength
        c = lambda t: t.to(dt)
        return RealRays(c(x0), c(y0), c(z0), c(L), c(M), c(N), c(torch.ones_like(x0)), c(w))

    # ---- RealRayTracer.trace / trace_generic, real_ray_tracer.py:37-133 -------------------------
    def trace(self, Hx, Hy, wavelength, num_rays=100, distribution="hexapolar", seed=None, spot=False):
        """N rays = fields x pupil points, field-major (real_ray_tracer.py:61-77). A call with several fields is ONE
        batch, like the reference's `trace_generic` with array-valued Hx, Hy: the Newton-Raphson exit test
        (newton_raphson.py:148) is a maximum over all its rays. Trace field by field (scalars, as SpotDiagram does) for
        per-field trip counts. `distribution`
        may be a name or a (Px, Py) pair of NumPy arrays / torch tensors (pinned host or device).
        spot=True also leaves the per-field spot sums (layout: include/olrt.h, olrt_spot_sums_*) in
        `rays.spot_sums` -- taken inside the trace kernel when the lens allows it (spot.spot_stats turns them
        into centroid / RMS radius)."""
        dev = be.get_device()
        if isinstance(distribution, str):
            if distribution == "random" and str(dev).startswith("cuda") and num_rays >= self.device_pupil_min_rays:
                Px, Py = random_pupil_device(num_rays, seed, dev)   # the seed travels, not the grid
            elif (distribution == "hexapolar" and str(dev).startswith("cuda")
                  and 1 + 3 * int(num_rays) * (int(num_rays) + 1) >= self.device_pupil_min_rays):
                Px, Py = hexapolar_pupil_device(num_rays, dev)      # index arithmetic on the device
            elif (distribution == "uniform" and str(dev).startswith("cuda")
                  and int(num_rays) * int(num_rays) >= self.device_pupil_min_rays):
                Px, Py = uniform_pupil_device(num_rays, dev)
            else:
                Px, Py = pupil_points(distribution, num_rays, seed)
        else:
            Px, Py = distribution
        Hx = np.atleast_1d(np.asarray(Hx, dtype=np.float64))
        Hy = np.atleast_1d(np.asarray(Hy, dtype=np.float64))
        sg = self.surface_group
        if (isinstance(Px, torch.Tensor) and not Px.is_cuda and Px.dtype == torch.float64
                and Px.numel() >= self.pipeline_min_rays and sg.record in ("none", "last", None)
                and not sg._wants_grad_params()):
            return self._trace_pipelined(Hx, Hy, Px, Py, wavelength, spot)
        if isinstance(Px, torch.Tensor):
            Px = Px.to(device=dev, dtype=torch.float64, non_blocking=True)
            Py = Py.to(device=dev, dtype=torch.float64, non_blocking=True)
        else:
            Px = torch.as_tensor(np.asarray(Px, dtype=np.float64), device=dev)
            Py = torch.as_tensor(np.asarray(Py, dtype=np.float64), device=dev)
        if (sg.fast_math and sg.record in ("none", "last", None) and not sg._wants_grad_params()
                and self._EPL is not None and self._EPD is not None and Px.numel() > 0):
            packed = sg.pack([float(wavelength)])
            if packed.fastkinds():
                # the grid is on the device: one fused launch + trace (+ spot) kernel over the whole pupil
                return self._trace_pipelined(Hx, Hy, Px.contiguous(), Py.contiguous(), wavelength, spot, packed)
        rays = self.launch_rays(Hx, Hy, Px, Py, wavelength)
        self.surface_group.trace(rays, wavelengths=[float(wavelength)])
        if spot:
            from .spot import spot_sums
            rays.spot_sums = spot_sums(rays.x, rays.y, rays.i, int(Hx.size), int(Px.numel()))
        return rays

    device_pupil_min_rays = 4096  # "random" distributions at least this long are generated on the device
    pipeline_min_rays = 1 << 21   # host pupil grids at least this long are streamed in chunks
    pipeline_chunks = 12
    pipeline_streamed = True     # fused kernel: one launch that waits for the chunks as they arrive
    pipeline_growth = 1.25       # chunk k + 1 is this much longer than chunk k

    def _trace_pipelined(self, Hx, Hy, Px_h, Py_h, wavelength, spot=False, packed=None):
        """Same result as launch_rays + SurfaceGroup.trace for a pupil grid that lives in (pinned) HOST
        memory: the grid is copied in chunks on a side stream while the chunks already on the device are
        traced, so the host->device copy overlaps the kernels instead of preceding them.

        Tolerance mode, lens of planes / conics / even aspheres: ONE kernel per chunk launches the rays of every
        field for the chunk's pupil points, traces them and (spot=True) accumulates the spot sums
        (`olrt_trace_pupil_*`); the launched rays never exist in memory. Otherwise each (chunk, field) slice is
        launched by `olrt_generate_rays_*` and traced by `olrt_trace_fwd_*`.

        The Newton trip counts stay batch-global: every chunk reports its residual maxima, ONE read-back checks
        them against the reference's exit rule (newton_raphson.py:148) and, if the counts were wrong for some
        chunk, the whole batch is traced again with the corrected counts."""
        from . import _lib as L
        from .trace import fused_trace, fused_trace_pupil, group_true_iters, slice_true_iters
        sg = self.surface_group
        dev = torch.device(be.get_device())
        dt = be.get_precision()
        nf, npup = int(Hx.size), int(Px_h.numel())
        n = nf * npup
        main = torch.cuda.current_stream(dev)
        if getattr(self, "_copy_stream", None) is None:
            self._copy_stream = torch.cuda.Stream(device=dev)
        cs = self._copy_stream
        resident = Px_h.is_cuda
        Px_d = Px_h if resident else torch.empty(npup, dtype=torch.float64, device=dev)
        Py_d = Py_h if resident else torch.empty(npup, dtype=torch.float64, device=dev)
        # pack first: it decides how the grid is streamed
        if packed is None:   # packed once, without records: the record mask does not decide eligibility, and in
            keep, sg.record = sg.record, "none"   # the fused path (record == "last") the final state IS the record
            try:
                packed = sg.pack([float(wavelength)])
            finally:
                sg.record = keep
            fused = bool(sg.fast_math and packed.fastkinds() and self._EPL is not None and self._EPD is not None)
            if not fused and keep != "none":
                packed = sg.pack([float(wavelength)])
        else:
            fused = bool(sg.fast_math and packed.fastkinds() and self._EPL is not None and self._EPD is not None)
            if fused and packed.record_surfaces:
                keep, sg.record = sg.record, "none"
                try:
                    packed = sg.pack([float(wavelength)])
                finally:
                    sg.record = keep
        # Streamed mode (fused kernel, host grid): ONE launch covers the whole grid while the copy engine is still
        # delivering it -- equal chunks, each followed on the copy stream by a 4-byte copy that sets the chunk's
        # "ready" word; the kernel's CTAs walk the pupil in copy order and wait on those words (olrt.h,
        # olrt_pupil.ready). Otherwise: one launch per chunk behind an event; the compute of chunk k must outlast
        # the copy of chunk k + 1 (the kernels are ~1.25x slower per pupil point than the PCIe copy), so the chunks
        # grow geometrically from a short first one -- the only copy nothing overlaps.
        C_ = 1 if resident else max(1, min(self.pipeline_chunks, npup // 65536))
        streamed = bool(fused and not resident and C_ > 1 and self.pipeline_streamed
                        and Px_h.is_pinned() and Py_h.is_pinned() and Px_h.is_contiguous() and Py_h.is_contiguous()
                        and Py_h.dtype == torch.float64)
        ready = None
        ready_chunk = 0
        if streamed:
            ready_chunk = (-(-npup // C_) + 255) // 256 * 256
            cuts = list(range(0, npup, ready_chunk)) + [npup]
            ready = torch.zeros(len(cuts) - 1, dtype=torch.int32, device=dev)
            if getattr(self, "_ones_pinned", None) is None or self._ones_pinned.numel() < len(cuts) - 1:
                self._ones_pinned = torch.ones(max(64, len(cuts) - 1), dtype=torch.int32).pin_memory()
        elif C_ >= 4:
            g_ = self.pipeline_growth
            w_ = np.cumsum([g_ ** k for k in range(C_)])
            cuts = [0] + [int(npup * v / w_[-1]) // 256 * 256 for v in w_[:-1]] + [npup]
        else:
            cuts = [c * npup // C_ for c in range(C_ + 1)]
        edges = [(a_, b_) for a_, b_ in zip(cuts[:-1], cuts[1:]) if b_ > a_]
        events = []
        tl = [] if getattr(self, "debug_timeline", False) else None   # [(label, timing event)]

        def mark(label, stream):
            if tl is not None:
                e = torch.cuda.Event(enable_timing=True)
                e.record(stream)
                tl.append((label, e))
        cs.wait_stream(main)        # (also orders the zeroing of `ready` before the copies that set it)
        mark("start", main)

        def enqueue_copies(upto):   # copies [len(events), upto) on the side stream, one event (and word) each
            with torch.cuda.stream(cs):
                for c_ in range(len(events), min(upto, len(edges))):
                    lo, hi = edges[c_]
                    if not resident:
                        Px_d[lo:hi].copy_(Px_h[lo:hi], non_blocking=True)
                        Py_d[lo:hi].copy_(Py_h[lo:hi], non_blocking=True)
                    if streamed:
                        ready[c_:c_ + 1].copy_(self._ones_pinned[c_:c_ + 1], non_blocking=True)
                    ev = torch.cuda.Event()
                    ev.record(cs)
                    events.append(ev)
                    mark(f"copy{c_}", cs)
        # streamed: ALL copies are queued before the kernel that waits for them is launched (it must never wait for
        # work the host has not submitted); otherwise the later ones are queued once the first kernel is on its way
        if streamed:
            L.check(L.load().olrt_upload_pupil(Px_d.data_ptr(), Py_d.data_ptr(), Px_h.data_ptr(), Py_h.data_ptr(), npup,
                                               ready_chunk, ready.data_ptr(), self._ones_pinned.data_ptr(),
                                               cs.cuda_stream), "olrt_upload_pupil")
            if tl is not None:
                mark("copies", cs)
        else:
            enqueue_copies(2)
        out = torch.empty((8, n), dtype=dt, device=dev)
        dir0 = torch.empty((3, n), dtype=dt, device=dev)
        ft = self._field_table(Hx, Hy).to(dev)
        fxy_all = ft.fxy_d
        gen_kw = dict(EPD=self._EPD, EPL=self._EPL, wavelength=float(wavelength), dtype=dt, **ft.kwargs())
        status_t = torch.zeros(1, dtype=torch.int32, device=dev)
        sums = shift = None
        if spot:  # sums are taken about each field's chief-ray image point (one tiny launch, no read-back)
            sums = torch.zeros((nf, L.SPOT_SUMS), dtype=torch.float64, device=dev)
            if fused:
                # the chief rays only CONDITION the second moments (the statistics do not depend on the point the sums
                # are taken about): traced once per lens state, fields and wavelength, then remembered on the lens
                ck_key = (packed.presc.tobytes(), packed.ops.tobytes(), ft.fxy.tobytes(), float(wavelength), str(dt),
                          str(dev))
                cache = self.__dict__.setdefault("_chief_shift_cache", {})
                shift = cache.get(ck_key)
                if shift is None:
                    zero = torch.zeros(1, dtype=torch.float64, device=dev)
                    chief = torch.empty((8, nf), dtype=dt, device=dev)
                    fused_trace_pupil(packed, zero, zero, fxy_all, out=chief, status_t=status_t, **gen_kw)
                    shift = chief[:2].t().to(torch.float64).nan_to_num(0.0, 0.0, 0.0).contiguous()
                    if len(cache) > 8:
                        cache.clear()
                    cache[ck_key] = shift
        launches = [(0, npup)] if streamed else edges
        rays_in = None if fused else torch.empty((8, n), dtype=dt, device=dev)
        n_slices = len(launches) * (1 if fused else nf)
        has_nr = bool(packed.nr_rows())
        # Newton trip counts PER FIELD in the fused path: the reference's Optic.trace takes one field per call, so each
        # field's rays decide their own exit test (newton_raphson.py:148); the kernel reads the counts of a tile's
        # field from `group_k` and reports per field (include/olrt.h, olrt_trace_pupil_*)
        per_field = bool(fused and has_nr and nf > 1)
        kf = None
        if per_field:   # first guess: what the same fields settled on last time (the exit rule is checked either way)
            kf_key = (tuple(np.asarray(Hx, dtype=np.float64).reshape(-1).tolist()),
                      tuple(np.asarray(Hy, dtype=np.float64).reshape(-1).tolist()), float(wavelength))
            kf_cache = sg.__dict__.setdefault("_field_counts_cache", {})
            kf = kf_cache.get(kf_key)
            if kf is None or kf.shape != (nf, packed.n_surfaces):
                kf = np.tile(np.asarray(packed.nr_iters(), dtype=np.int32), (nf, 1))
            kf = kf.copy()
        for attempt in range(8):
            if per_field:
                packed.set_nr_iters(kf.max(axis=0).tolist())   # the slot layout covers every field's counts
                gk_cache = sg.__dict__.setdefault("_group_k_device", {})    # the counts rarely change between calls
                gk_key = (kf.tobytes(), kf.shape, str(dev))
                group_k = gk_cache.get(gk_key)
                if group_k is None:
                    if len(gk_cache) > 16:
                        gk_cache.clear()
                    group_k = gk_cache[gk_key] = torch.from_numpy(kf.copy()).to(dev)
                resid = torch.zeros((nf, packed.n_surfaces, L.GROUP_RESID), dtype=torch.int64, device=dev)
            else:
                group_k = None
                resid = (torch.zeros((n_slices, packed.n_surfaces, L.RESID_STRIDE), dtype=torch.int64, device=dev)
                         if has_nr else None)
            k_used = packed.nr_iters()
            if attempt and sums is not None:
                sums.zero_()
            q_ = 0
            for c_, (lo, hi) in enumerate(launches):
                if c_ == 1 and not streamed:
                    enqueue_copies(len(edges))
                if not streamed:
                    main.wait_event(events[c_])
                if fused:
                    fused_trace_pupil(packed, Px_d, Py_d, fxy_all, p_lo=lo, p_len=hi - lo, out=out, dir0=dir0,
                                      spot=sums, spot_shift=shift, status_t=status_t,
                                      resid_out=None if (resid is None or per_field) else resid[q_],
                                      group_k=group_k, group_resid=resid if per_field else None, ready=ready,
                                      ready_chunk=ready_chunk, **gen_kw)
                    mark(f"chunk{q_}", main)
                    q_ += 1
                    continue
                for f in range(nf):
                    a, b = f * npup + lo, f * npup + hi
                    self.launch_rays(Hx[f:f + 1], Hy[f:f + 1], Px_d[lo:hi], Py_d[lo:hi], wavelength,
                                     out=rays_in[:, a:b], table=ft.rows(f))
                    fused_trace(packed, [rays_in[q, a:b] for q in range(7)] + [None], out=out[:, a:b],
                                dir0_out=dir0[:, a:b], fast=sg.fast_math, verify_nr=False, check_status=False,
                                status_t=status_t, resid_out=None if resid is None else resid[q_])
                    mark(f"slice{q_}", main)
                    q_ += 1
            # ONE read-back for the whole batch (pinned staging, one wait): the residual maxima (the batch-global
            # exit rule is the max over the slices), the status word and, in the fused path, the spot sums
            if spot and fused:
                sums[:, 8:10] = shift
            pins = self._pinned_tail(resid, status_t, sums if (spot and fused) else None)
            done = torch.cuda.Event()
            done.record(main)
            done.synchronize()
            st = int(pins[1][0])
            sums_host = None if pins[2] is None else pins[2].clone()
            if not has_nr:
                break
            if per_field:
                kt = group_true_iters(packed, pins[0].numpy().view(np.float64), kf)
                if (kt == kf).all():
                    break
                if ((kt < 0) & (kf >= L.GROUP_RESID - 1)).any():
                    raise L.OlrtLibraryError("a field needs more than %d Newton updates per surface"
                                             % (L.GROUP_RESID - 1))
                kf = np.where(kt < 0, np.minimum(kf * 2 + 1, L.GROUP_RESID - 1), kt).astype(np.int32)
                continue
            rows = pins[0].numpy().view(np.float64)
            per_slice = [slice_true_iters(packed, rows[q], k_used) for q in range(n_slices)]
            if any(v is None for ps in per_slice for v in ps):      # some slice needs more updates
                packed.set_nr_iters([min(lay.max_iter, max(k * 2, k + 2)) if lay.geom >= 2 else 0
                                     for k, lay in zip(k_used, packed.layouts)])
                continue
            batch = [max(col) for col in zip(*per_slice)]
            if batch == k_used:
                break
            packed.set_nr_iters(batch)
        if streamed:
            main.wait_stream(cs)    # the copies are done (the kernel consumed them); this tells the allocator so
        if st & L.ST_TIMEOUT:
            raise L.OlrtLibraryError("streamed pupil grid: a host->device chunk never arrived (OLRT_ST_TIMEOUT)")
        if tl is not None:
            torch.cuda.synchronize(dev)
            self.last_timeline = [(lb, tl[0][1].elapsed_time(e)) for lb, e in tl]
        if st & 1:
            raise ValueError("Zernike coordinates must be normalized to [-1, 1]. Consider updating "
                             "the normalization radius to 1.1x the surface aperture.")
        if fused:
            rays = RealRays.from_traced(out, dir0, wavelength)
        else:
            rays = RealRays(*[rays_in[q] for q in range(8)])
        sg.reset()
        sg._finish(rays, packed, out, None, dir0, packed.nr_iters())
        rays.nr_iters_per_field = kf.tolist() if per_field else [packed.nr_iters()] * nf
        if per_field:
            if len(kf_cache) > 16:
                kf_cache.clear()
            kf_cache[kf_key] = kf.copy()
        if fused and sg.record == "last":
            sg.surfaces[-1]._set_record(out)
        if spot:
            if fused:
                rays.spot_sums_host = sums_host   # the same sums, already on the host (read back with the status)
            else:
                from .spot import spot_sums
                spot_sums(rays.x, rays.y, rays.i, nf, npup, out=sums)
            rays.spot_sums = sums
        return rays

    def _pinned_tail(self, resid, status_t, sums):
        """Queue the device->host copies of the small per-call results into cached pinned buffers (no sync here).
        Returns (resid_host or None, status_host, sums_host or None)."""
        cache = self.__dict__.setdefault("_pin_cache", {})
        outs = []
        for name, t in (("resid", resid), ("status", status_t), ("sums", sums)):
            if t is None:
                outs.append(None)
                continue
            key = (name, t.dtype, tuple(t.shape))
            h = cache.get(key)
            if h is None:
                h = torch.empty(t.shape, dtype=t.dtype).pin_memory()
                cache[key] = h
            h.copy_(t, non_blocking=True)
            outs.append(h)
        return outs

    def _field_offsets(self, Hx, Hy):
        """(fxy [F, 2] NumPy, z_start): -tan(radians(max_field H)) (offset + EPL) per field and the launch plane,
        field_types.py:139-169, :223-239 (object at infinity)."""
        EPL, EPD, mf = self._EPL, self._EPD, self.max_field
        z_pos = [float(s.geometry.cs.z) for s in self.surface_group.surfaces[1:-1]]
        offset = EPD - min(z_pos)
        hx, hy = np.asarray(Hx, dtype=np.float64).reshape(-1), np.asarray(Hy, dtype=np.float64).reshape(-1)
        fxy = np.stack([-np.tan(np.radians(mf * hx)) * (offset + EPL),
                        -np.tan(np.radians(mf * hy)) * (offset + EPL)], axis=1)
        return np.ascontiguousarray(fxy), float(self.surface_group.surfaces[1].geometry.cs.z) - offset

    def _object_sag(self, x, y):
        """Sag of the object surface at (x, y) (NumPy, fp64): 0 for a plane, geometries/standard.py:73-87 for a
        conic -- what ObjectHeightField.get_ray_origins adds to the object's z (field_types.py:271)."""
        g = self.surface_group.surfaces[0].geometry
        name = type(g).__name__
        if name == "Plane":
            return np.zeros_like(np.asarray(x, dtype=np.float64))
        if name == "StandardGeometry":
            R, k = float(g.radius), float(g.k)
            r2 = x ** 2 + y ** 2
            return r2 / (R * (1 + np.sqrt(1 - (1 + k) * r2 / R ** 2)))
        raise NotImplementedError(f"object-height fields on a {name} object surface")

    def _field_table(self, Hx, Hy) -> FieldTable:
        """The per-field launch scalars for fields (Hx[f], Hy[f]) (host, NumPy fp64, the reference's expressions):
        AngleField.get_ray_origins field_types.py:139-169, ObjectHeightField.get_ray_origins :245-272,
        vx = 1 - vig_x ray_generator.py:50-52."""
        hx, hy = np.asarray(Hx, dtype=np.float64).reshape(-1), np.asarray(Hy, dtype=np.float64).reshape(-1)
        vig = None
        if any(v != (0.0, 0.0) for v in self.vignetting):
            vxf, vyf = self.get_vig_factor(hx, hy)
            vig = np.ascontiguousarray(np.stack([1 - vxf, 1 - vyf], axis=1))
        obj = self.surface_group.surfaces[0]
        mf = self.max_field
        if self.field_type == "object_height":
            if obj.is_infinite:
                raise ValueError('Field type cannot be "object_height" for an object at infinity.')
            x0, y0 = mf * hx, mf * hy
            fz = self._object_sag(x0, y0) + float(obj.geometry.cs.z)
            return FieldTable(np.ascontiguousarray(np.stack([x0, y0], axis=1)), vig, np.ascontiguousarray(fz), 0.0)
        if obj.is_infinite:
            fxy, z_start = self._field_offsets(hx, hy)
            return FieldTable(fxy, vig, None, z_start)
        zo = float(obj.geometry.cs.z)
        EPL = self._EPL
        fxy = np.stack([-np.tan(np.radians(mf * hx)) * (EPL - zo), -np.tan(np.radians(mf * hy)) * (EPL - zo)], axis=1)
        return FieldTable(np.ascontiguousarray(fxy), vig, np.full(hx.size, zo), 0.0)

    def launch_rays(self, Hx, Hy, Px, Py, wavelength, out=None, table=None) -> RealRays:
        """fields x pupil rays (field-major) from the ray-launch kernel `olrt_generate_rays_*`
        (ray_generator.py:28-106 + field_types.py:139-169, :245-272). Hx, Hy: [F] NumPy arrays
        (one entry per field); Px, Py: [P] float64 device tensors. `table`: a FieldTable already on the device
        (a caller launching many slices uploads it once and passes `table.rows(f)`)."""
        import ctypes as C

        from . import _lib as L
        if self._EPL is None or self._EPD is None:
            raise ValueError("call set_paraxial(EPL, EPD) first")
        lib = L.load()
        dt, dev = be.get_precision(), Px.device
        EPL, EPD = self._EPL, self._EPD
        if table is None:
            table = self._field_table(Hx, Hy).to(dev)
        nf, npup = table.n_fields, int(Px.numel())
        if out is None:
            out = torch.empty((8, nf * npup), dtype=dt, device=dev)
        fn = lib.olrt_generate_rays_f64 if dt == torch.float64 else lib.olrt_generate_rays_f32
        Px, Py = Px.contiguous(), Py.contiguous()
        from .trace import _ptr
        pu = L.OlrtPupil(Px.data_ptr(), Py.data_ptr(), table.fxy_d.data_ptr(), _ptr(table.vig_d), _ptr(table.fz_d),
                         npup, 0, npup, nf, 0, float(EPD), float(EPL), table.z_start, float(wavelength))
        with torch.cuda.device(dev):
            rc = fn(C.byref(pu), L.PtrArray8(*[out[q].data_ptr() for q in range(8)]),
                    torch.cuda.current_stream(dev).cuda_stream)
        L.check(rc, "olrt_generate_rays")
        if not out.is_contiguous():
            return None  # a slice of a larger batch: the caller owns the RealRays object
        return RealRays(*[out[q] for q in range(8)])

    def trace_generic(self, Hx, Hy, Px, Py, wavelength):
        arrs = [np.atleast_1d(np.asarray(a, dtype=np.float64)) for a in (Hx, Hy, Px, Py)]
        n = max(a.size for a in arrs)
        arrs = [np.full(n, a[0]) if a.size == 1 else a for a in arrs]
        rays = self.generate_rays(*arrs, wavelength)
        self.surface_group.trace(rays, wavelengths=[float(wavelength)])
        return rays
