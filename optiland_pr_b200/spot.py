"""Spot epilogue: per-(field, wavelength) centroid / RMS / geometric radius from device-side sums.

    RayOperand.rms_spot_size                 optiland/optimization/operand/ray.py:299-340
    SpotDiagram.centroid / rms_spot_radius / geometric_spot_radius
                                             optiland/analysis/spot_diagram.py:317-357

The reductions run in `olrt_spot_sums_*` (csrc/olrt_api.cu: registers -> warp shuffles -> one set
of atomics per CTA); this module only turns the ten sums per group into the scalars the
reference computes with `be.mean` / `be.sqrt` / `be.max` over whole arrays.
"""
from __future__ import annotations

import torch

from . import _lib as L


def spot_sums(x: torch.Tensor, y: torch.Tensor, intensity: torch.Tensor, n_groups: int = 1,
              rays_per_group: int | None = None, shift: torch.Tensor | None = None,
              valid_only: bool = False, out: torch.Tensor | None = None) -> torch.Tensor:
    """[n_groups, 10] float64 device tensor, layout in include/olrt.h (olrt_spot_sums_*).
    Ray r belongs to group r // rays_per_group. Asynchronous on the current stream."""
    lib = L.load()
    if not x.is_cuda:
        raise L.OlrtLibraryError("spot_sums needs CUDA tensors; there is no CPU path")
    n = int(x.numel())
    if rays_per_group is None:
        rays_per_group = (n + n_groups - 1) // n_groups if n else 1
    x, y, intensity = x.contiguous(), y.contiguous(), intensity.contiguous()
    if out is None:
        out = torch.zeros((n_groups, L.SPOT_SUMS), dtype=torch.float64, device=x.device)
    else:
        out.zero_()
    if shift is not None:
        shift = shift.to(device=x.device, dtype=torch.float64).contiguous()
    fn = lib.olrt_spot_sums_f64 if x.dtype == torch.float64 else lib.olrt_spot_sums_f32
    with torch.cuda.device(x.device):
        rc = fn(x.data_ptr(), y.data_ptr(), intensity.data_ptr(), n, int(rays_per_group), int(n_groups),
                0 if shift is None else shift.data_ptr(), L.SPOT_VALID_ONLY if valid_only else 0,
                out.data_ptr(), torch.cuda.current_stream(x.device).cuda_stream)
    L.check(rc, "olrt_spot_sums")
    return out


def spot_stats(sums: torch.Tensor) -> dict:
    """count, centroid and RMS radius about the centroid per group, from one pass of sums taken
    about any shift: mean((x-cx)^2 + (y-cy)^2) = (Sxx + Syy)/n - mx^2 - my^2 with mx = Sx/n."""
    cnt = sums[:, 0]
    mx, my = sums[:, 1] / cnt, sums[:, 2] / cnt
    var = (sums[:, 3] + sums[:, 4]) / cnt - mx * mx - my * my
    return dict(count=cnt, cx=sums[:, 8] + mx, cy=sums[:, 9] + my,
                rms=torch.sqrt(torch.clamp(var, min=0.0)), intensity=sums[:, 6])


class _RmsSpotSq(torch.autograd.Function):
    """mean((x - mean x)^2 + (y - mean y)^2) per group of a traced batch `out` [8, N] (field-major, equal groups), all
    rays counted as RayOperand.rms_spot_size does (operand/ray.py:336-340). Forward: two passes of the spot-sum kernel
    (centroid, then sums about it). Backward: d / d x_i = 2 (x_i - mean x) / n exactly (the centroid's own dependence
    cancels) -- ONE dense [8, N] gradient with two filled rows instead of the ~30 elementwise kernels autograd records
    for the same expression."""

    @staticmethod
    def forward(ctx, out, n_groups):
        x, y, inten = out[0], out[1], out[6]
        n = int(x.numel())
        per = n // n_groups
        if per * n_groups != n:
            raise ValueError("rms_spot_sq needs equal groups")
        st = spot_stats(spot_sums(x, y, inten, n_groups, per))   # sums about each group's first ray + that shift
        cen = torch.stack([st["cx"], st["cy"]], dim=1).contiguous()
        s2 = spot_sums(x, y, inten, n_groups, per, shift=cen)
        cnt = s2[:, 0]
        val = (s2[:, 3] + s2[:, 4]) / cnt - (s2[:, 1] / cnt) ** 2 - (s2[:, 2] / cnt) ** 2
        ctx.save_for_backward(out, cen)
        ctx.shape = (n_groups, per)
        return val

    @staticmethod
    def backward(ctx, g):
        out, cen = ctx.saved_tensors
        ng, per = ctx.shape
        G = torch.zeros_like(out)
        scale = (g * (2.0 / per)).to(out.dtype).unsqueeze(1)
        c = cen.to(out.dtype)
        torch.mul(out[0].view(ng, per) - c[:, 0:1], scale, out=G[0].view(ng, per))
        torch.mul(out[1].view(ng, per) - c[:, 1:2], scale, out=G[1].view(ng, per))
        return G, None


def rms_spot_sq(out: torch.Tensor, n_groups: int) -> torch.Tensor:
    """Squared RMS spot radius per group, differentiable w.r.t. the traced batch `out` [8, N] (see _RmsSpotSq)."""
    return _RmsSpotSq.apply(out, n_groups)


def rms_spot_size(x, y, intensity, n_groups=1, rays_per_group=None, valid_only=False):
    """sqrt(mean((x - mean x)^2 + (y - mean y)^2)) per group (ray.py:336-340 when valid_only is
    False: vignetted and NaN rays are included, exactly as the reference's operand does)."""
    return spot_stats(spot_sums(x, y, intensity, n_groups, rays_per_group, None, valid_only))["rms"]


def spot_radii(x, y, intensity, n_groups=1, rays_per_group=None, reference_centroid=None):
    """SpotDiagram-style numbers (valid rays only): centroid, RMS radius and geometric radius about
    `reference_centroid` ([G, 2]; default: each group's own centroid). Two passes."""
    s1 = spot_sums(x, y, intensity, n_groups, rays_per_group, None, True)
    st = spot_stats(s1)
    c = reference_centroid if reference_centroid is not None else torch.stack([st["cx"], st["cy"]], dim=1)
    s2 = spot_sums(x, y, intensity, n_groups, rays_per_group, c, True)
    rms = torch.sqrt((s2[:, 3] + s2[:, 4]) / s2[:, 0])
    geo = torch.sqrt(s2[:, 5])  # stored as the bits of a double (u64-ordered atomicMax)
    return dict(cx=c[:, 0], cy=c[:, 1], rms=rms, geo=geo, count=s2[:, 0])


class SpotDiagram:
    """analysis/spot_diagram.py:203-357 without the plotting: centroid(), rms_spot_radius() and
    geometric_spot_radius() for fields x wavelengths.

    The reference traces one (field, wavelength) pair per Python call (spot_diagram.py:387-395) and keeps every ray;
    here ALL pairs go through one fused launch + trace + spot kernel per pass: the wavelengths are
    rows of a prescription batch (the media of a row are the indices at its wavelength), the fields share the pupil
    grid, the launched and traced rays never exist in memory. Pass 1 takes the sums about the chief-ray image points
    (-> centroids of the reference wavelength, spot_diagram.py:317-327); pass 2 takes them about those centroids
    (-> RMS and maximum radius, :329-357). Rays with intensity <= 0 or NaN are left out, as :427-429 does.
    Lenses outside the fused kernel's surface kinds go through Optic.trace per pair and the two-pass `spot_radii`."""

    def __init__(self, optic, fields="all", wavelengths="all", num_rings=6, distribution="hexapolar"):
        import numpy as np

        from . import backend as be
        from .optic import pupil_points
        self.optic = optic
        self.fields = optic.get_field_coords() if fields == "all" else list(fields)
        self.wavelengths = list(optic.wavelengths) if wavelengths == "all" else [float(w) for w in wavelengths]
        ref_w = optic.primary_wavelength
        self.ref_index = self.wavelengths.index(ref_w) if ref_w in self.wavelengths else 0
        dev = torch.device(be.get_device())
        dt = be.get_precision()
        if isinstance(distribution, str):
            px, py = pupil_points(distribution, num_rings)
        else:
            px, py = distribution
        Px = torch.as_tensor(px, dtype=torch.float64, device=dev).contiguous()
        Py = torch.as_tensor(py, dtype=torch.float64, device=dev).contiguous()
        nf, nw = len(self.fields), len(self.wavelengths)
        Hx = np.array([f[0] for f in self.fields], dtype=np.float64)
        Hy = np.array([f[1] for f in self.fields], dtype=np.float64)
        sg = optic.surface_group
        keep = sg.record
        sg.record = "none"
        try:
            # ONE pack; the rows of the other wavelengths differ from it in the media entries only
            from .prescription import media_row
            p0 = sg.pack([self.wavelengths[0]])
            rows = [p0.presc[0]] + [media_row(p0, sg.surfaces, w) for w in self.wavelengths[1:]]
            if any(r is None for r in rows):       # a medium absorbs at some wavelengths only: full packs
                packs = [p0] + [sg.pack([w]) for w in self.wavelengths[1:]]
                rows = [p.presc[0] for p in packs] if all(np.array_equal(p.ops, p0.ops) for p in packs) else None
        finally:
            sg.record = keep
        fused = bool(sg.fast_math and p0.fastkinds() and optic._EPL is not None and optic._EPD is not None
                     and rows is not None)
        s2 = None
        if fused:
            try:
                s2, cen = self._fused_sums(optic, p0, np.stack(rows), Px, Py, Hx, Hy, dev, dt, nf, nw)
            except NotImplementedError:
                # a (field, wavelength) group needs more Newton updates than the fused launch holds -- e.g. a ray
                # that misses an asphere makes the reference run max_iter of them (newton_raphson.py:137-149):
                # trace pair by pair, every trace with its own counts
                fused = False
        self.path = "fused" if fused else "per_pair"
        if not fused:
            npup = int(Px.numel())
            xs, ys, is_ = [], [], []
            for w in self.wavelengths:      # one trace per pair: the Newton exit rule is per trace
                for f in range(nf):
                    rays = optic.trace(Hx[f:f + 1], Hy[f:f + 1], w, distribution=(Px, Py))
                    xs.append(rays.x), ys.append(rays.y), is_.append(rays.i)
            x, y, i = torch.cat(xs), torch.cat(ys), torch.cat(is_)
            s1 = spot_sums(x, y, i, nw * nf, npup, None, True)
            st = spot_stats(s1)
            cen = torch.stack([st["cx"], st["cy"]], dim=1).reshape(nw, nf, 2)[self.ref_index]
            shift2 = cen.unsqueeze(0).expand(nw, nf, 2).reshape(-1, 2).contiguous()
            s2 = spot_sums(x, y, i, nw * nf, npup, shift2, True)
        # a handful of numbers per (field, wavelength) pair: ONE read-back, the arithmetic on the host (each tiny
        # device op would cost more in launch overhead than the whole lot)
        s2 = s2.cpu()
        self._centroid = cen.cpu()
        self._rms = torch.sqrt((s2[:, 3] + s2[:, 4]) / s2[:, 0]).reshape(nw, nf)
        self._geo = torch.sqrt(s2[:, 5]).reshape(nw, nf)
        self.count = s2[:, 0].reshape(nw, nf)

    def _fused_sums(self, optic, p0, rows, Px, Py, Hx, Hy, dev, dt, nf, nw):
        """Both passes through trace.fused_spot_groups; returns (second-pass sums [nw * nf, 10], centroids [nf, 2])."""
        import numpy as np

        from .trace import fused_spot_groups
        ft = optic._field_table(Hx, Hy).to(dev)
        gen_kw = dict(EPD=optic._EPD, EPL=optic._EPL, wavelength=float(self.wavelengths[0]), dtype=dt, **ft.kwargs())
        rows = np.array(rows, dtype=np.float64)      # (a copy: the scaling below must not reach p0's own row)
        # the kernels form the Beer-Lambert exponent 4 pi k / lambda (homogeneous.py:49) with the launch's ONE
        # wavelength (the first row's): rows of other wavelengths carry k lambda_0 / lambda_row instead of k
        for r, w in enumerate(self.wavelengths):
            if r:
                for lay in p0.layouts:
                    rows[r, lay.media + 2] *= float(self.wavelengths[0]) / float(w)
        presc = torch.from_numpy(rows).to(dev)
        k0 = np.zeros((nf, p0.n_surfaces), dtype=np.int32)
        for k in p0.nr_rows():
            k0[:, k] = 2
        # ONE trace: the first pass also keeps the final ray states (64 B per ray) when they fit comfortably; the second
        # pass -- sums about the reference wavelength's centroids -- is then a streaming reduction of the kept x, y, i
        # (olrt_spot_sums_*), not a second trace
        npup = int(Px.numel())
        keep = None
        if nw * nf * npup * 64 <= self.keep_bytes_limit:
            keep = torch.empty((8, nw * nf * npup), dtype=dt, device=dev)
        s1, kg = fused_spot_groups(p0, presc, Px, Py, ft.fxy_d, gen_kw, k0, valid_only=True, keep=keep)
        st = spot_stats(s1.cpu())     # (the launch above has been waited for: its Newton reports were read)
        cen = torch.stack([st["cx"], st["cy"]], dim=1).reshape(nw, nf, 2)[self.ref_index]     # [nf, 2], host
        shift2 = cen.unsqueeze(0).expand(nw, nf, 2).reshape(-1, 2).contiguous().to(dev)
        if keep is not None and getattr(keep, "valid", False):
            s2 = spot_sums(keep[0], keep[1], keep[6], nw * nf, npup, shift2, True)
            self.passes = 1
        else:
            s2, _ = fused_spot_groups(p0, presc, Px, Py, ft.fxy_d, gen_kw, kg, shift=shift2, valid_only=True,
                                      settle=False)
            self.passes = 2
        self.group_iters = kg
        return s2, cen

    keep_bytes_limit = 8 << 30     # final states kept for the second pass up to this many bytes, else two traces

    def centroid(self):
        """[(cx, cy)] per field, reference wavelength (spot_diagram.py:317-327)."""
        c = self._centroid.cpu()
        return [(c[f, 0], c[f, 1]) for f in range(c.shape[0])]

    def rms_spot_radius(self):
        """[field][wavelength] (spot_diagram.py:344-357)."""
        r = self._rms.cpu()
        return [[r[w, f] for w in range(r.shape[0])] for f in range(r.shape[1])]

    def geometric_spot_radius(self):
        """[field][wavelength] (spot_diagram.py:329-342)."""
        r = self._geo.cpu()
        return [[r[w, f] for w in range(r.shape[0])] for f in range(r.shape[1])]
