"""Optic facade for the hot path: the calls analyses and operands make before reading
`surface_group.x[-1, :]`.

    Optic.add_surface / add_field / add_wavelength / set_aperture   optiland/optic/optic.py
    Optic.trace(Hx, Hy, wavelength, num_rays, distribution)         optic.py:584-609
    Optic.trace_generic(Hx, Hy, Px, Py, wavelength)                 optic.py:611-632
    RealRayTracer.trace / trace_generic                             raytrace/real_ray_tracer.py:37-133
    RayGenerator.generate_rays (infinite-object AngleField branch)  rays/ray_generator.py:28-106,
                                                                    fields/field_types.py:139-169

Scope notes: ray generation here is the infinite-conjugate angle-field case
without vignetting factors; the first-order quantities EPL / EPD it needs are INPUTS
(`set_paraxial`) because the paraxial tracer (optiland/paraxial.py) is outside the hot path.
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


class Optic:
    def __init__(self, name=None):
        self.name = name
        self.surface_group = SurfaceGroup()
        self.fields = []          # (x, y) in degrees
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
        if field_type != "angle":
            raise ValueError("only angle fields are implemented")

    def add_field(self, y=0.0, x=0.0):
        self.fields.append((float(x), float(y)))

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
        return max(math.hypot(fx, fy) for fx, fy in self.fields) if self.fields else 0.0

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

        Hx, Hy, Px, Py = f64(Hx), f64(Hy), f64(Px), f64(Py)
        EPL, EPD, mf = self._EPL, self._EPD, self.max_field
        z_pos = [float(s.geometry.cs.z) for s in self.surface_group.surfaces[1:-1]]
        offset = EPD - min(z_pos)                              # field_types.py:223-239
        x = -torch.tan(torch.deg2rad(mf * Hx)) * (offset + EPL)
        y = -torch.tan(torch.deg2rad(mf * Hy)) * (offset + EPL)
        z = float(self.surface_group.surfaces[1].geometry.cs.z) - offset
        x0 = Px * EPD / 2 * 1.0 + x
        y0 = Py * EPD / 2 * 1.0 + y
        z0 = torch.full_like(x0, z)
        x1 = Px * EPD * 1.0 / 2
        y1 = Py * EPD * 1.0 / 2
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
        """N rays = fields x pupil points, field-major (real_ray_tracer.py:61-77). `distribution`
        may be a name or a (Px, Py) pair of NumPy arrays / torch tensors (pinned host or device).
        spot=True also leaves the per-field spot sums (layout: include/olrt.h, olrt_spot_sums_*) in
        `rays.spot_sums` -- taken inside the trace kernel when the lens allows it (spot.spot_stats turns them
        into centroid / RMS radius)."""
        if isinstance(distribution, str):
            Px, Py = pupil_points(distribution, num_rays, seed)
        else:
            Px, Py = distribution
        dev = be.get_device()
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
                and self._EPL is not None and self._EPD is not None and Px.numel() > 0
                and sg.pack([float(wavelength)]).fastkinds()):
            # the grid is on the device: one fused launch + trace (+ spot) kernel over the whole pupil
            return self._trace_pipelined(Hx, Hy, Px.contiguous(), Py.contiguous(), wavelength, spot)
        rays = self.launch_rays(Hx, Hy, Px, Py, wavelength)
        self.surface_group.trace(rays, wavelengths=[float(wavelength)])
        if spot:
            from .spot import spot_sums
            rays.spot_sums = spot_sums(rays.x, rays.y, rays.i, int(Hx.size), int(Px.numel()))
        return rays

    pipeline_min_rays = 1 << 21   # host pupil grids at least this long are streamed in chunks
    pipeline_chunks = 4

    def _trace_pipelined(self, Hx, Hy, Px_h, Py_h, wavelength, spot=False):
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
        from .trace import fused_trace, fused_trace_pupil, slice_true_iters
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
        out = torch.empty((8, n), dtype=dt, device=dev)
        dir0 = torch.empty((3, n), dtype=dt, device=dev)
        # chunk edges: a short first chunk (the only copy nothing overlaps), then equal ones
        C_ = 1 if resident else max(1, min(self.pipeline_chunks, npup // 65536))
        if C_ >= 4:
            cuts = [0, npup // (2 * C_), npup // C_] + [c * npup // C_ for c in range(2, C_ + 1)]
        else:
            cuts = [c * npup // C_ for c in range(C_ + 1)]
        edges = [(a_, b_) for a_, b_ in zip(cuts[:-1], cuts[1:]) if b_ > a_]
        fxy_np, z_start = self._field_offsets(Hx, Hy)
        fxy_all = torch.from_numpy(fxy_np).to(dev, non_blocking=True)
        packed = sg.pack([float(wavelength)])
        fused = bool(sg.fast_math and packed.fastkinds() and self._EPL is not None and self._EPD is not None)
        if fused and packed.record_surfaces:   # record == "last": the final state IS the last surface's record
            keep, sg.record = sg.record, "none"
            try:
                packed = sg.pack([float(wavelength)])
            finally:
                sg.record = keep
        gen_kw = dict(EPD=self._EPD, EPL=self._EPL, z_start=z_start, wavelength=float(wavelength), dtype=dt)
        status_t = torch.zeros(1, dtype=torch.int32, device=dev)
        sums = shift = None
        if spot:  # sums are taken about each field's chief-ray image point (one tiny launch, no read-back)
            sums = torch.zeros((nf, L.SPOT_SUMS), dtype=torch.float64, device=dev)
            if fused:
                zero = torch.zeros(1, dtype=torch.float64, device=dev)
                chief = torch.empty((8, nf), dtype=dt, device=dev)
                fused_trace_pupil(packed, zero, zero, fxy_all, out=chief, status_t=status_t, **gen_kw)
                shift = chief[:2].t().to(torch.float64).nan_to_num(0.0, 0.0, 0.0).contiguous()
        cs.wait_stream(main)
        events = []
        tl = [] if getattr(self, "debug_timeline", False) else None   # [(label, timing event)]

        def mark(label, stream):
            if tl is not None:
                e = torch.cuda.Event(enable_timing=True)
                e.record(stream)
                tl.append((label, e))
        mark("start", main)
        with torch.cuda.stream(cs):
            for c_, (lo, hi) in enumerate(edges):
                if not resident:
                    Px_d[lo:hi].copy_(Px_h[lo:hi], non_blocking=True)
                    Py_d[lo:hi].copy_(Py_h[lo:hi], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(cs)
                events.append(ev)
                mark(f"copy{c_}", cs)
        rays_in = None if fused else torch.empty((8, n), dtype=dt, device=dev)
        n_slices = len(edges) * (1 if fused else nf)
        has_nr = bool(packed.nr_rows())
        for attempt in range(8):
            k_used = packed.nr_iters()
            resid = (torch.zeros((n_slices, packed.n_surfaces, L.RESID_STRIDE), dtype=torch.int64, device=dev)
                     if has_nr else None)
            if attempt and sums is not None:
                sums.zero_()
            q_ = 0
            for (lo, hi), ev in zip(edges, events):
                main.wait_event(ev)
                if fused:
                    fused_trace_pupil(packed, Px_d, Py_d, fxy_all, p_lo=lo, p_len=hi - lo, out=out, dir0=dir0,
                                      spot=sums, spot_shift=shift, status_t=status_t,
                                      resid_out=None if resid is None else resid[q_], **gen_kw)
                    mark(f"chunk{q_}", main)
                    q_ += 1
                    continue
                for f in range(nf):
                    a, b = f * npup + lo, f * npup + hi
                    self.launch_rays(Hx[f:f + 1], Hy[f:f + 1], Px_d[lo:hi], Py_d[lo:hi], wavelength,
                                     out=rays_in[:, a:b], fxy_d=fxy_all[f:f + 1])
                    fused_trace(packed, [rays_in[q, a:b] for q in range(7)] + [None], out=out[:, a:b],
                                dir0_out=dir0[:, a:b], fast=sg.fast_math, verify_nr=False, check_status=False,
                                status_t=status_t, resid_out=None if resid is None else resid[q_])
                    mark(f"slice{q_}", main)
                    q_ += 1
            if not has_nr:
                break
            # ONE read-back for the whole batch: the batch-global exit rule is the max over the slices
            rows = resid.cpu().numpy().view(np.float64)
            per_slice = [slice_true_iters(packed, rows[q], k_used) for q in range(n_slices)]
            if any(v is None for ps in per_slice for v in ps):      # some slice needs more updates
                packed.set_nr_iters([min(lay.max_iter, max(k * 2, k + 2)) if lay.geom >= 2 else 0
                                     for k, lay in zip(k_used, packed.layouts)])
                continue
            batch = [max(col) for col in zip(*per_slice)]
            if batch == k_used:
                break
            packed.set_nr_iters(batch)
        st = int(status_t.item())
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
        if fused and sg.record == "last":
            sg.surfaces[-1]._set_record(out)
        if spot:
            if fused:
                sums[:, 8:10] = shift
            else:
                from .spot import spot_sums
                spot_sums(rays.x, rays.y, rays.i, nf, npup, out=sums)
            rays.spot_sums = sums
        return rays

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

    def launch_rays(self, Hx, Hy, Px, Py, wavelength, out=None, fxy_d=None) -> RealRays:
        """fields x pupil rays (field-major) from the ray-launch kernel `olrt_generate_rays_*`
        (ray_generator.py:28-106 + field_types.py:139-169, object at infinity). Hx, Hy: [F] NumPy arrays
        (one entry per field); Px, Py: [P] float64 device tensors."""
        import ctypes as C

        from . import _lib as L
        if self._EPL is None or self._EPD is None:
            raise ValueError("call set_paraxial(EPL, EPD) first")
        lib = L.load()
        dt, dev = be.get_precision(), Px.device
        EPL, EPD = self._EPL, self._EPD
        fxy, z_start = self._field_offsets(Hx, Hy)
        if fxy_d is None:      # a caller launching many slices uploads the table once and passes row views
            fxy_d = torch.from_numpy(fxy).to(dev, non_blocking=True)
        nf, npup = int(fxy.shape[0]), int(Px.numel())
        if out is None:
            out = torch.empty((8, nf * npup), dtype=dt, device=dev)
        fn = lib.olrt_generate_rays_f64 if dt == torch.float64 else lib.olrt_generate_rays_f32
        Px, Py = Px.contiguous(), Py.contiguous()
        with torch.cuda.device(dev):
            rc = fn(Px.data_ptr(), Py.data_ptr(), npup, fxy_d.data_ptr(), nf, float(EPD), float(EPL), z_start,
                    float(wavelength), L.PtrArray8(*[out[q].data_ptr() for q in range(8)]),
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
