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
perturbations outside that kernel (tilts) go through launch + trace + spot kernels on chunks of trials.

Sharding: trials are independent; rank r takes a contiguous range of them and the per-trial merits are gathered
(`parallel.py`), nothing else crosses NVLink.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib as L
from . import backend as be
from . import parallel
from .trace import fused_trace, fused_trace_pupil, slice_true_iters


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
        """Fill `vals` [rows, n_perturbations] with the next draws of every sampler, trial by trial."""
        for t in range(vals.shape[0]):           # monte_carlo.py:86-92: per trial, every perturbation in order
            for j, p in enumerate(self.perturbations):
                p.value = p.sampler.sample()
                vals[t, j] = p.value

    def draw(self, num_iterations):
        vals = np.empty((num_iterations, len(self.perturbations)))
        self._draw_rows(vals)
        self.values = vals
        return vals

    draw_chunks = 4      # fused path, >= 1024 trials: the draws of chunk c + 1 run on the host while chunk c traces

    def prescriptions(self, packed, values):
        """[n_trials, stride] float64: the nominal row with every trial's perturbed entries."""
        base = np.ascontiguousarray(packed.presc[0], dtype=np.float64)
        batch = np.repeat(base[None, :], values.shape[0], axis=0)
        for j, p in enumerate(self.perturbations):
            if p.variable_type == "tilt":
                raise NotImplementedError("tilt perturbations change the ops table: use run(), not prescriptions()")
            for off, scale, nominal in _offsets(packed, self.optic, p):
                batch[:, off] += scale * (values[:, j] - nominal)
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
        fused = bool(packed.fastkinds())
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
            batch = self.prescriptions(packed, values)
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
        self.last_path = "fused" if fused else "chunked"
        step = trials_per_launch or (hi - lo if fused else max(1, (1 << 26) // max(1, nf * npup)))
        zero1 = torch.zeros(1, dtype=torch.float64, device=dev)
        has_nr = bool(packed.nr_rows())
        from .spot import spot_stats, spot_sums
        from .trace import fused_spot_groups
        S = packed.n_surfaces
        nr_rows = packed.nr_rows()

        def fused_groups(presc_t, k0, between=None):
            """(rms [T, nf], counts [T * nf, S]) from trace.fused_spot_groups: one fused launch per round over
            `presc_t` [T, stride], Newton trip counts per (trial, field) group."""
            sums, kg = fused_spot_groups(packed, presc_t, Px, Py, fxy, gen_kw, k0, between=between)
            return spot_stats(sums)["rms"].reshape(int(presc_t.shape[0]), nf), kg

        k_nominal = np.zeros((nf, S), dtype=np.int32)
        for k in nr_rows:
            k_nominal[:, k] = 2
        if fused:
            # the nominal lens first: its per-field counts are the guess for every trial
            nominal = torch.from_numpy(np.ascontiguousarray(packed.presc[:1], dtype=np.float64)).to(dev)
            _, kg0 = fused_groups(nominal, k_nominal)
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
                    self.group_iters = kg if self.group_iters is None else np.concatenate([self.group_iters, kg])
                elif nxt is not None:
                    nxt()
            lo_loop = hi       # nothing left for the generic loop below
        else:
            lo_loop = lo
        for a in range(lo_loop, hi, max(1, step)):
            b = min(hi, a + step)
            presc = torch.from_numpy(batch[a:b]).to(dev)
            T = b - a
            if fused:
                rms_ab, kg = fused_groups(presc, k_nominal)
                rms_local[a - lo:b - lo] = rms_ab
                self.group_iters = kg if self.group_iters is None else np.concatenate([self.group_iters, kg])
                continue
            for attempt in range(8):
                status = torch.zeros(1, dtype=torch.int32, device=dev)
                resid = (torch.zeros((S, L.RESID_STRIDE), dtype=torch.int64, device=dev) if has_nr else None)
                sums = torch.zeros((T * nf, L.SPOT_SUMS), dtype=torch.float64, device=dev)
                rays = lens.launch_rays(Hx, Hy, Px, Py, wavelength)           # one copy of the launched rays
                ins = [t.repeat(T) for t in rays.components()[:7]] + [None]   # trial-major batch
                res = fused_trace(packed, ins, presc=presc, rays_per_presc=nf * npup, fast=sg.fast_math,
                                  verify_nr=False, check_status=False, status_t=status, resid_out=resid)
                spot_sums(res.out[0], res.out[1], res.out[6], T * nf, npup, out=sums)
                if not has_nr:
                    break
                k_used = packed.nr_iters()
                k_true = slice_true_iters(packed, resid.cpu().numpy().view(np.float64), k_used)
                if k_true == k_used:
                    break
                packed.set_nr_iters([min(lay.max_iter, max(k * 2, k + 2)) if kt is None else kt
                                     for k, kt, lay in zip(k_used, k_true, packed.layouts)])
            rms_local[a - lo:b - lo] = spot_stats(sums)["rms"].reshape(T, nf)
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
