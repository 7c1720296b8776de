"""Monte-Carlo tolerancing on batched prescriptions (BASELINE config C5).

    DistributionSampler / RangeSampler / Perturbation   optiland/tolerancing/perturbation.py:41-180
    MonteCarlo.run                                       optiland/tolerancing/monte_carlo.py:60-123
    RayOperand.rms_spot_size (the merit)                 optiland/optimization/operand/ray.py:299-340

The reference re-traces the lens trial by trial in Python: reset, apply every perturbation (one scalar draw per
perturbation per trial, each perturbation owning its own seeded generator), evaluate the operands. Here the draws
are made in the same order (so the perturbation values are the reference's, bit for bit), every trial becomes one
row of a `[n_trials, stride]` prescription batch, and ONE kernel launch traces all trials: each trial launches the
same pupil grid for every field, traces it through its own prescription and reduces its spot sums in the kernel
(`olrt_trace_pupil_*` with olrt_sys.n_presc > 1) -- no per-ray data exists in HBM at any point. Lenses or
perturbations outside that kernel (tilts, freeform surfaces), and trials whose Newton counts the fused launch cannot
hold (a missing ray makes the reference run max_iter updates), go trial by trial through launch + trace + spot
kernels -- one launch per trial, each with its own Newton counts, exactly like the reference's loop.

Sharding: trials are independent; rank r takes a contiguous range of them and the per-trial merits are gathered
(`parallel.py`), nothing else crosses NVLink.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib as L
from . import backend as be
from . import parallel
from .trace import fused_trace


class RangeSampler:
    """perturbation.py:41-88: cycles through linspace(start, end, steps)."""

    def __init__(self, start, end, steps):
        self.values = np.linspace(start, end, steps)
        self.index = 0
        self.size = steps

    def sample(self):
        if self.index >= len(self.values):
            self.index = 0
        v = self.values[self.index]
        self.index += 1
        return float(v)

    def sample_n(self, n):
        """The next n values of `sample()` at once."""
        idx = (self.index % len(self.values) + np.arange(n)) % len(self.values)
        self.index = int(idx[-1]) + 1 if n else self.index
        return self.values[idx].astype(np.float64)


class DistributionSampler:
    """perturbation.py:91-134: one scalar draw per call from numpy's default_rng(seed)."""

    def __init__(self, distribution, seed=None, **params):
        self.generator = np.random.default_rng(seed)
        self.distribution = distribution
        self.params = params
        for p in params:
            if p not in ("loc", "scale", "low", "high"):
                raise ValueError(f"Invalid parameter: {p}")

    def sample(self):
        if self.distribution == "normal":
            return float(self.generator.normal(**self.params))
        if self.distribution == "uniform":
            return float(self.generator.uniform(**self.params))
        raise ValueError(f"Unknown distribution: {self.distribution}")

    def sample_n(self, n):
        """The next n values of `sample()` in one call: numpy's Generator fills an array with the same routine it
        uses for a scalar, so the values (and the generator's state afterwards) are those of n scalar draws."""
        if self.distribution == "normal":
            return self.generator.normal(size=n, **self.params)
        if self.distribution == "uniform":
            return self.generator.uniform(size=n, **self.params)
        raise ValueError(f"Unknown distribution: {self.distribution}")


class Perturbation:
    """perturbation.py:137-180: `variable_type` in radius / conic / thickness / decenter (axis "x" | "y") /
    asphere_coeff (coeff_number) / tilt (axis "x" | "y", radians); the sampled value REPLACES the variable
    (samplers are centred on the nominal value, as in the reference)."""

    def __init__(self, optic, variable_type, sampler, surface_number, axis=None, coeff_number=None):
        self.optic, self.variable_type, self.sampler = optic, variable_type, sampler
        self.surface_number, self.axis, self.coeff_number = surface_number, axis, coeff_number
        self.value = None

    def __str__(self):
        extra = self.axis if self.axis is not None else (self.coeff_number if self.coeff_number is not None else "")
        return f"{self.variable_type}{extra}, Surface {self.surface_number}"


def _offsets(packed, lens, pert):
    """[(offset in the prescription row, scale, nominal)]: new_entry = entry + scale * (value - nominal)."""
    surfaces = lens.surface_group.surfaces
    by_index = {lay.index: lay for lay in packed.layouts}
    lay = by_index.get(pert.surface_number)
    g = surfaces[pert.surface_number].geometry
    vt = pert.variable_type
    if vt == "thickness":   # optic_updater.set_thickness: every later surface moves along z by the change
        nominal = float(surfaces[pert.surface_number].thickness)
        return [(by_index[k].frames + 2, 1.0, nominal) for k in by_index if k > pert.surface_number]
    if lay is None:
        raise ValueError("surface is not traced")
    if vt == "radius":
        return [(lay.radius, 1.0, float(g.radius))]
    if vt == "conic":
        return [(lay.conic, 1.0, float(g.k))]
    if vt == "decenter":
        q = {"x": 0, "y": 1}[pert.axis]
        return [(lay.frames + q, 1.0, float(getattr(g.cs, pert.axis)))]
    if vt == "asphere_coeff":
        return [(lay.coef + pert.coeff_number, 1.0, float(g.coefficients[pert.coeff_number]))]
    raise NotImplementedError(vt)


class MonteCarloSpot:
    """RMS spot size per (trial, field) for `num_iterations` perturbed copies of `optic`."""

    def __init__(self, optic, perturbations):
        self.optic, self.perturbations = optic, list(perturbations)
        self.values = None      # [n_trials, n_perturbations], the draws (trial-major, the reference's order)
        self.last_path = None

    def _draw_rows(self, vals):
        """Fill `vals` [rows, n_perturbations] with the next draws of every sampler. The reference draws trial by
        trial, every perturbation in order (monte_carlo.py:86-92) -- but every sampler owns its generator
        (perturbation.py:91-134), so the t-th value of a perturbation is its sampler's t-th draw whatever the
        interleaving: one vectorised draw per perturbation gives the same values. Samplers without `sample_n`
        (a caller's own) are drawn one by one."""
        T = vals.shape[0]
        for j, p in enumerate(self.perturbations):
            if hasattr(p.sampler, "sample_n"):
                vals[:, j] = p.sampler.sample_n(T)
            else:
                for t in range(T):
                    vals[t, j] = p.sampler.sample()
            if T:
                p.value = float(vals[-1, j])

    def draw(self, num_iterations):
        vals = np.empty((num_iterations, len(self.perturbations)))
        self._draw_rows(vals)
        self.values = vals
        return vals

    draw_chunks = 4      # fused path, >= 1024 trials: the draws of chunk c + 1 run on the host while chunk c traces

    def has_tilt(self) -> bool:
        return any(p.variable_type == "tilt" for p in self.perturbations)

    def prescriptions(self, packed, values, ops_out=None):
        """[n_trials, stride] float64: the nominal row with every trial's perturbed entries. Tilt perturbations
        (cs.rx / ry / rz := value, radians) also decide a frame flag of the ops table -- the reference applies a
        rotation only when the angle is truthy (coordinate_system.py:73-107) -- so they need `ops_out`, a list that
        receives one int32 ops table per trial."""
        base = np.ascontiguousarray(packed.presc[0], dtype=np.float64)
        batch = np.repeat(base[None, :], values.shape[0], axis=0)
        tilts = []
        for j, p in enumerate(self.perturbations):
            if p.variable_type == "tilt":
                if ops_out is None:
                    raise NotImplementedError("tilt perturbations change the ops table: pass ops_out (run() does)")
                tilts.append((j, p))
                continue
            for off, scale, nominal in _offsets(packed, self.optic, p):
                batch[:, off] += scale * (values[:, j] - nominal)
        if ops_out is not None:
            rows = {lay.index: k for k, lay in enumerate(packed.layouts)}
            for t in range(values.shape[0]):
                ops = packed.ops.copy()
                for j, p in tilts:
                    k = rows[p.surface_number]
                    lay = packed.layouts[k]
                    q = lay.nframes - 1                      # the surface's own frame is the last of its chain
                    c_off, s_off, bit = {"z": (3, 4, 1), "y": (5, 6, 2), "x": (7, 8, 4)}[p.axis]
                    v = np.float64(values[t, j])
                    batch[t, lay.frames + q * L.FRAME_REALS + c_off] = np.cos(v)
                    batch[t, lay.frames + q * L.FRAME_REALS + s_off] = np.sin(v)
                    word = int(ops[k * L.OPS_STRIDE + L.OP_FRAMEFLAGS]) & ~(bit << (3 * q))
                    if bool(v):
                        word |= bit << (3 * q)
                    ops[k * L.OPS_STRIDE + L.OP_FRAMEFLAGS] = word
                ops_out.append(ops)
        return batch

    def run(self, num_iterations, Hx, Hy, wavelength, Px, Py, trials_per_launch=None):
        """rms [n_trials, n_fields] (float64, on every rank). Px, Py: pupil grid (NumPy or tensors)."""
        lens, sg = self.optic, self.optic.surface_group
        dev = torch.device(be.get_device())
        dt = be.get_precision()
        keep = sg.record
        sg.record = "none"
        try:
            packed = sg.pack([float(wavelength)])
        finally:
            sg.record = keep
        fused = bool(packed.fastkinds()) and not self.has_tilt()
        # Fused path: the reference-order scalar draws (one Python call per trial and perturbation) are made chunk by
        # chunk, each chunk while the previous one is on the GPU; every sampler keeps its own generator, so the values
        # are the ones one pass over all trials would draw. Every rank draws every chunk (same generator state).
        chunked_draws = fused and num_iterations >= 1024 and self.draw_chunks > 1 and trials_per_launch is None
        if chunked_draws:
            values = np.empty((num_iterations, len(self.perturbations)))
            self.values = values
            draw_edges = [int(v) for v in np.linspace(0, num_iterations, self.draw_chunks + 1)]
            self._draw_rows(values[draw_edges[0]:draw_edges[1]])
            batch = None
        else:
            values = self.draw(num_iterations)
            ops_rows = [] if self.has_tilt() else None
            batch = self.prescriptions(packed, values, ops_out=ops_rows)
        lo, hi = parallel.shard_range(num_iterations)
        Hx = np.atleast_1d(np.asarray(Hx, dtype=np.float64))
        Hy = np.atleast_1d(np.asarray(Hy, dtype=np.float64))
        nf = int(Hx.size)
        Px = torch.as_tensor(np.asarray(Px, dtype=np.float64) if not isinstance(Px, torch.Tensor) else Px,
                             dtype=torch.float64, device=dev).contiguous()
        Py = torch.as_tensor(np.asarray(Py, dtype=np.float64) if not isinstance(Py, torch.Tensor) else Py,
                             dtype=torch.float64, device=dev).contiguous()
        npup = int(Px.numel())
        ft = lens._field_table(Hx, Hy).to(dev)
        fxy = ft.fxy_d
        gen_kw = dict(EPD=lens._EPD, EPL=lens._EPL, wavelength=float(wavelength), dtype=dt, **ft.kwargs())
        rms_local = torch.empty((hi - lo, nf), dtype=torch.float64, device=dev)
        self.last_path = "fused" if fused else "per_trial"
        self.fallback_trials = 0
        step = trials_per_launch or (hi - lo)
        from .spot import spot_stats, spot_sums
        from .trace import fused_spot_groups
        S = packed.n_surfaces
        nr_rows = packed.nr_rows()

        launched = []      # the launched rays of the nominal fields, made once for the per-trial path

        def per_trial(rows, ops_list=None):
            """rms [T, nf] trial by trial, the reference's own loop shape (monte_carlo.py:86-109): one launch per
            trial through `fused_trace`, which settles that trial's Newton counts with the reference's exit rule
            (a NaN ray -> max_iter). Any lens the kernels trace, any ray count, tilts included."""
            import dataclasses
            if not launched:
                rays = lens.launch_rays(Hx, Hy, Px, Py, wavelength)
                launched.append(rays.components()[:7] + [None])
            out_rms = torch.empty((rows.shape[0], nf), dtype=torch.float64, device=dev)
            for t in range(rows.shape[0]):
                pk = dataclasses.replace(packed, presc=np.ascontiguousarray(rows[t:t + 1]),
                                         ops=(ops_list[t] if ops_list is not None else packed.ops).copy())
                res = fused_trace(pk, launched[0], fast=sg.fast_math)
                packed.set_nr_iters(res.nr_iters)             # the next trial's first guess
                if res.status & L.ST_ZERNIKE_RANGE:
                    raise ValueError("Zernike coordinates must be normalized to [-1, 1]. Consider updating the "
                                     "normalization radius to 1.1x the surface aperture.")
                out_rms[t] = spot_stats(spot_sums(res.out[0], res.out[1], res.out[6], nf, npup))["rms"]
            return out_rms

        def fused_groups(presc_t, k0, between=None, rows=None):
            """(rms [T, nf], counts [T * nf, S]) from trace.fused_spot_groups: one fused launch per round over
            `presc_t` [T, stride], Newton trip counts per (trial, field) group. Where a group needs more Newton
            updates than the fused launch holds (OLRT_GROUP_RESID - 1; a ray that misses an asphere makes the
            reference run max_iter of them) the batch goes trial by trial instead."""
            try:
                sums, kg = fused_spot_groups(packed, presc_t, Px, Py, fxy, gen_kw, k0, between=between)
            except NotImplementedError:
                self.fallback_trials += int(presc_t.shape[0])
                rows_np = rows if rows is not None else presc_t.cpu().numpy()
                return per_trial(rows_np), None
            return spot_stats(sums)["rms"].reshape(int(presc_t.shape[0]), nf), kg

        k_nominal = np.zeros((nf, S), dtype=np.int32)
        for k in nr_rows:
            k_nominal[:, k] = 2
        if fused:
            # the nominal lens first: its per-field counts are the guess for every trial
            nominal = torch.from_numpy(np.ascontiguousarray(packed.presc[:1], dtype=np.float64)).to(dev)
            _, kg0 = fused_groups(nominal, k_nominal)
            if kg0 is not None:
                k_nominal = kg0.reshape(nf, S)
        self.group_iters = None
        if chunked_draws:
            for c in range(self.draw_chunks):
                a, b = max(draw_edges[c], lo), min(draw_edges[c + 1], hi)
                nxt = None
                if c + 1 < self.draw_chunks:
                    nxt = (lambda c_=c: self._draw_rows(values[draw_edges[c_ + 1]:draw_edges[c_ + 2]]))
                if b > a:
                    presc = torch.from_numpy(self.prescriptions(packed, values[a:b])).to(dev)
                    rms_ab, kg = fused_groups(presc, k_nominal, between=nxt)
                    rms_local[a - lo:b - lo] = rms_ab
                    if kg is not None:
                        self.group_iters = kg if self.group_iters is None else np.concatenate([self.group_iters, kg])
                elif nxt is not None:
                    nxt()
            lo_loop = hi       # nothing left for the generic loop below
        else:
            lo_loop = lo
        for a in range(lo_loop, hi, max(1, step)):
            b = min(hi, a + step)
            if fused:
                presc = torch.from_numpy(batch[a:b]).to(dev)
                rms_ab, kg = fused_groups(presc, k_nominal, rows=batch[a:b])
                rms_local[a - lo:b - lo] = rms_ab
                if kg is not None:
                    self.group_iters = kg if self.group_iters is None else np.concatenate([self.group_iters, kg])
                continue
            rms_local[a - lo:b - lo] = per_trial(batch[a:b], ops_rows[a:b] if ops_rows is not None else None)
        r, w = parallel.world()
        if w == 1:
            return rms_local
        import torch.distributed as dist
        sizes = [parallel.shard_range(num_iterations, q, w) for q in range(w)]
        pad = max(h - l for l, h in sizes)
        buf = torch.zeros((pad, nf), dtype=torch.float64, device=dev)
        buf[:hi - lo] = rms_local
        parts = [torch.empty_like(buf) for _ in range(w)]
        dist.all_gather(parts, buf)
        return torch.cat([parts[q][:h - l] for q, (l, h) in enumerate(sizes)])
