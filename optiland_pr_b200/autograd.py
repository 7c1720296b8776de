"""torch.autograd.Function over the fused forward / backward kernels.

What it replaces: torch autograd recording every elementwise op of `Surface.trace`
(optiland/surfaces/standard_surface.py:186-233) when the reference runs on its torch backend with
`be.grad_mode.enable()` (optimization/optimizer/torch/base.py:95-154 calls loss.backward() through
that graph). Here the graph has ONE node: forward = `olrt_trace_fwd_*` recording every surface (the
records are the checkpoints), backward = `olrt_trace_bwd_*`.

The prescription enters the node as one flat tensor assembled by `prescription_tensor` from the live
parameter tensors (geometry.radius / .k / .coefficients[i], cs.x / .y / .z), so gradients flow on to
whatever those depend on (e.g. cs.z of later surfaces on an earlier thickness, optic_updater.py:67-85).
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np
import torch

from . import _lib as L
from .prescription import PackedSystem
from .trace import _device_tables, _fixup_workspace, _ptr, fused_trace


def _live_entries(packed: PackedSystem, surfaces):
    """[(offset in the prescription, tensor)] for every parameter that is a torch tensor."""
    out = []
    for lay in packed.layouts:
        surf = surfaces[lay.index]
        g = surf.geometry
        cs = g.cs
        for q, nm in enumerate(("x", "y", "z")):
            v = getattr(cs, nm)
            if isinstance(v, torch.Tensor):
                out.append((lay.frames + q, v))
        if lay.geom != L.GEOM_PLANE:
            if isinstance(g.radius, torch.Tensor):
                out.append((lay.radius, g.radius))
            if isinstance(getattr(g, "k", None), torch.Tensor):
                out.append((lay.conic, g.k))
        if lay.geom in (L.GEOM_EVEN_ASPHERE, L.GEOM_ODD_ASPHERE):
            for i, c in enumerate(g.coefficients):
                if isinstance(c, torch.Tensor):
                    out.append((lay.coef + i, c))
        if lay.geom == L.GEOM_POLYNOMIAL:      # C[i][j] row-major (geometries/polynomial.py:93-111)
            co = g.coefficients
            for i in range(len(co)):
                for j in range(len(co[i])):
                    c = co[i][j]
                    if isinstance(c, torch.Tensor):
                        out.append((lay.coef + i * len(co[i]) + j, c))
        if lay.geom == L.GEOM_ZERNIKE:         # per term the row holds c and c N (optimization/variable/zernike_coeff.py:71)
            co = g.zernike.coeffs
            for j, (_, _, off, norm) in enumerate(lay.zernike_terms):
                c = co[j]
                if isinstance(c, torch.Tensor):
                    out.append((off, c))
                    out.append((off + 1, c * norm))
        # tilts (optimization/variable/tilt.py:59): the row holds cos / sin of rz, ry, rx of the surface's one frame
        if getattr(cs, "reference_cs", None) is None:
            for q, nm in enumerate(("rz", "ry", "rx")):
                v = getattr(cs, nm, None)
                if isinstance(v, torch.Tensor):
                    out.append((lay.frames + 3 + 2 * q, torch.cos(v)))
                    out.append((lay.frames + 4 + 2 * q, torch.sin(v)))
    return out


def prescription_tensor(packed: PackedSystem, surfaces, dtype, device) -> torch.Tensor:
    """Flat prescription [stride] on `device`, differentiable w.r.t. the live parameter tensors."""
    base = torch.from_numpy(np.ascontiguousarray(packed.presc[0])).to(device=device, dtype=dtype)
    live = _live_entries(packed, surfaces)
    # offsets that can receive a gradient: the backward skips the frame x / y columns when none of them is live
    packed.live_offsets = frozenset(o for o, t in live if t.requires_grad)
    if not live:
        return base
    idx = torch.tensor([o for o, _ in live], dtype=torch.int64, device=device)
    vals = torch.stack([t.to(device=device, dtype=dtype).reshape(()) for _, t in live])
    return base.scatter(0, idx, vals)


# OLRT_CHECK_BWD_STATUS=1: also read the backward kernel's status word back after every launch (a device sync)
_CHECK_BWD_STATUS = os.environ.get("OLRT_CHECK_BWD_STATUS", "0") == "1"


class FusedTraceFn(torch.autograd.Function):
    """(presc, x, y, z, L, M, N, i, opd_in, w) -> (out [8, N], record [S, 8, N])."""

    @staticmethod
    def forward(ctx, presc, x, y, z, Lc, M, N, inten, opd_in, w, packed, fast, generic_adjoint=False,
                blocked=False):
        if not blocked and packed.record_surfaces != [lay.index for lay in packed.layouts]:
            raise L.OlrtLibraryError("the differentiable trace needs every surface recorded (record='all')")
        if blocked and (packed.record_surfaces or not fast or generic_adjoint or not packed.reduced_adjoint()):
            raise NotImplementedError("blocked checkpoints: record='none', tolerance mode, plane / conic / "
                                      "even-asphere lens, reduced-algebra adjoint")
        ins = [x, y, z, Lc, M, N, inten, w]
        res = fused_trace(packed, [t.detach() if t is not None else None for t in ins],
                          opd_in=None if opd_in is None else opd_in.detach(),
                          presc=presc.detach().reshape(1, -1), want_dir0=True, fast=fast, alias_out=True,
                          ckpt=blocked, own_rows=True,   # the row was assembled from this very pack
                          # a lens of planes / conics / even aspheres can only raise the packer-bug bits of the status
                          # word (no Zernike / Chebyshev range errors): do not stall the optimiser step on reading it
                          # or the fix-up count back (OLRT_CHECK_BWD_STATUS=1 reads both)
                          check_status=_CHECK_BWD_STATUS or not packed.fastkinds(), read_fixups=_CHECK_BWD_STATUS)
        if blocked and max(res.nr_iters, default=0) > L.BX_MAXK:
            raise NotImplementedError("blocked checkpoints: a surface took more than 3 Newton updates")
        ctx.blocked = bool(blocked)
        ctx.ckpt = res.ckpt
        live = getattr(packed, "live_offsets", None)     # set by prescription_tensor; None: a caller's own leaf row
        ctx.no_frame_xy = live is not None and not any(o in live for lay in packed.layouts
                                                       for o in (lay.frames, lay.frames + 1))
        ctx.set_materialize_grads(False)
        ctx.packed = packed
        ctx.generic_adjoint = bool(generic_adjoint)
        ctx.nr_iters = list(res.nr_iters)
        ctx.save_for_backward(presc.detach(), *[t.detach() for t in ins if t is not None],
                              *([] if blocked else [res.record]))
        ctx.has_w = w is not None
        ctx.has_opd = opd_in is not None
        ctx.status = res.status
        ctx.dir0 = res.dir0
        if blocked:   # no [S, 8, N] records in this mode: the second output is an empty placeholder
            empty = res.out.new_empty((0, 8, 0))
            ctx.mark_non_differentiable(res.dir0, empty)
            return res.out, empty, res.dir0
        ctx.mark_non_differentiable(res.dir0)
        return res.out, res.record, res.dir0

    @staticmethod
    def backward(ctx, gout, grecord, _gdir0):
        packed = ctx.packed
        saved = ctx.saved_tensors
        if ctx.blocked:
            presc, record, grecord = saved[0], ctx.ckpt, None
            ins = list(saved[1:])
        else:
            presc, record = saved[0], saved[-1]
            ins = list(saved[1:-1])
        if not ctx.has_w:
            ins.append(None)
        dtype, device = ins[0].dtype, ins[0].device
        n = int(ins[0].numel())
        lib = L.load()
        if not packed.differentiable(require_records=not ctx.blocked):
            raise NotImplementedError(
                "olrt_trace_bwd: gradients are implemented for plane / conic / even- or odd-asphere / x^i y^j "
                "polynomial / Zernike surfaces on one (decentred, tilted) frame; this lens has a surface outside "
                "that set")
        packed.set_nr_iters(ctx.nr_iters)
        ops_d, _ = _device_tables(packed, dtype, device)
        # d loss / d rays only where a ray tensor asks for it (the kernel writes every element of a requested array)
        need = [bool(ctx.needs_input_grad[1 + q]) for q in range(7)] + [bool(ctx.has_opd and ctx.needs_input_grad[8])]
        gin = torch.empty((8, n), dtype=dtype, device=device) if any(need) else None
        gpresc = torch.zeros_like(presc)
        status = torch.zeros(1, dtype=torch.int32, device=device)
        gout = None if gout is None else gout.contiguous()
        grecord = None if grecord is None else grecord.contiguous()
        sys_ = packed.make_sys()
        sys_.presc_stride = int(presc.numel())
        fn = lib.olrt_trace_bwd_f64 if dtype == torch.float64 else lib.olrt_trace_bwd_f32
        in_arr = L.PtrArray8(*[_ptr(t) for t in ins])
        gout_arr = L.PtrArray8(*[0 if gout is None else gout[q].data_ptr() for q in range(8)])
        gin_arr = L.PtrArray8(*[gin[q].data_ptr() if need[q] else 0 for q in range(8)])
        # the reduced-algebra adjoint (csrc/trace_bx.cuh) keeps <= BX_MAXK Newton iterates in registers
        flags = L.BWD_GENERIC if (ctx.generic_adjoint or max(ctx.nr_iters, default=0) > L.BX_MAXK
                                  or not packed.reduced_adjoint()) else 0
        if ctx.blocked:
            if flags:
                raise NotImplementedError("blocked checkpoints: a surface took more than 3 Newton updates; trace "
                                          "with record='all'")
            flags = L.BWD_CKPT_BLOCKED
        if ctx.no_frame_xy:
            flags |= L.BWD_NO_FRAME_XY
        with torch.cuda.device(device):
            rc = fn(C.byref(sys_), presc.data_ptr(), ops_d.data_ptr(), n, in_arr, record.data_ptr(), gout_arr,
                    _ptr(grecord), gin_arr, gpresc.data_ptr(), status.data_ptr(), flags,
                    torch.cuda.current_stream(device).cuda_stream)
        if rc == -4:
            raise NotImplementedError("olrt_trace_bwd: this lens is outside the differentiable set")
        L.check(rc, "olrt_trace_bwd")
        # the host check above mirrors the kernels' own; reading the device word back would serialise every step
        if _CHECK_BWD_STATUS and int(status.item()) & L.ST_BAD_HINT:
            raise NotImplementedError(
                "olrt_trace_bwd: the kernel found a surface outside the differentiable set (plane / conic / asphere / "
                "polynomial / Zernike on one frame, include/olrt.h)")
        g = [gin[q] if need[q] else None for q in range(8)]
        return (gpresc, g[0], g[1], g[2], g[3], g[4], g[5], g[6], g[7], None, None, None, None, None)


def trace_differentiable(packed: PackedSystem, presc: torch.Tensor, rays_in, opd_in=None, fast: bool = False,
                         generic_adjoint: bool = False, blocked: bool = False):
    """rays_in: 8 tensors (w may be None for a single wavelength). Returns (out, record, dir0).
    generic_adjoint=True forces the general backward kernel (csrc/trace_bwd.cuh) instead of the reduced-algebra one.
    blocked=True (packed with record="none", tolerance mode, plane / conic / even-asphere lens): the checkpoints are
    tile-blocked (include/olrt.h, OLRT_FWD_CKPT_BLOCKED) -- 48 B per ray and surface, read by the backward at fixed
    offsets from one pointer -- and `record` comes back empty."""
    x, y, z, Lc, M, N, inten, w = (list(rays_in) + [None])[:8]
    return FusedTraceFn.apply(presc, x, y, z, Lc, M, N, inten, opd_in, w, packed, fast, generic_adjoint, blocked)


def _rms_merit(lens, fields, wavelength, Px, Py):
    """sum over fields of rms_spot_size^2 (RayOperand.rms_spot_size, operand/ray.py:299-340 fed to
    OptimizationProblem.sum_squared, problem.py:100-109) from ONE traced batch of fields x pupil rays."""
    nf = len(fields)
    rays = lens.trace([f[0] for f in fields], [f[1] for f in fields], wavelength, distribution=(Px, Py))
    x = rays.x.reshape(nf, -1)
    y = rays.y.reshape(nf, -1)
    r2 = (x - x.mean(dim=1, keepdim=True)) ** 2 + (y - y.mean(dim=1, keepdim=True)) ** 2
    return r2.mean(dim=1).sum()


def _rms_merit_fused(lens, fields, wavelength, Px, Py):
    """The same merit with the spot reduction and its gradient in the library's own epilogue (spot.rms_spot_sq: two
    passes of the spot-sum kernel forward, one dense gradient backward) instead of recorded torch elementwise ops."""
    from .spot import rms_spot_sq
    nf = len(fields)
    rays = lens.trace([f[0] for f in fields], [f[1] for f in fields], wavelength, distribution=(Px, Py))
    lt = getattr(lens.surface_group, "last_trace", None)
    if lt is not None and lt.out is not None and lt.out[0].data_ptr() == rays.x.data_ptr():
        out = lt.out          # the [8, N] output of this very trace (rays.x ... are its rows)
    else:
        out = torch.stack([rays.x, rays.y, rays.z, rays.L, rays.M, rays.N, rays.i, rays.opd])
    return rms_spot_sq(out, nf).sum()


def _rms_merit_all(lens, fields, wavelengths, Px, Py):
    """sum over wavelengths x fields of rms_spot_size^2 from ONE traced batch: the rays of every (wavelength, field)
    pair carry their wavelength (the kernels index the media per ray), so an optimisation step is one FusedTraceFn
    node -- one pack, one launch pair, one backward -- instead of one per wavelength. Px, Py: device tensors [P]."""
    from . import backend as be
    from .spot import rms_spot_sq
    nf, nw, npup = len(fields), len(wavelengths), int(Px.numel())
    dev = Px.device
    Hx = torch.tensor([f[0] for f in fields], dtype=torch.float64, device=dev).repeat_interleave(npup).repeat(nw)
    Hy = torch.tensor([f[1] for f in fields], dtype=torch.float64, device=dev).repeat_interleave(npup).repeat(nw)
    wl = torch.tensor([float(w) for w in wavelengths], dtype=torch.float64, device=dev).repeat_interleave(nf * npup)
    rays = lens.generate_rays(Hx, Hy, Px.repeat(nf * nw), Py.repeat(nf * nw), wl)
    sg = lens.surface_group
    sg.trace(rays, wavelengths=[float(w) for w in wavelengths])
    lt = sg.last_trace
    if lt is not None and lt.out is not None and lt.out[0].data_ptr() == rays.x.data_ptr():
        out = lt.out
    else:
        out = torch.stack([rays.x, rays.y, rays.z, rays.L, rays.M, rays.N, rays.i, rays.opd])
    return rms_spot_sq(out, nw * nf).sum()


def _rms_merit_pairs(lens, pairs, Px, Py):
    """sum of rms_spot_size^2 over the given ((Hx, Hy), wavelength) pairs from ONE traced batch (per-ray wavelengths:
    one FusedTraceFn node). `pairs` may be any subset of fields x wavelengths -- a rank's share of them."""
    from . import backend as be
    from .rays import RealRays
    from .spot import rms_spot_sq
    npair, npup = len(pairs), int(Px.numel())
    dev = Px.device
    # the rays of every pair from the ray-launch kernel (one launch per wavelength: its fields share the pupil grid),
    # pair after pair in one [8, npair * npup] batch -- no elementwise torch ops, nothing recorded by autograd (the
    # launch depends on EPL / EPD only, which this package takes as inputs)
    wls = sorted({float(p[1]) for p in pairs})
    buf = torch.empty((8, npair * npup), dtype=be.get_precision(), device=dev)
    a = 0
    for w in wls:
        fw = [p[0] for p in pairs if float(p[1]) == w]
        b = a + len(fw) * npup
        lens.launch_rays(np.array([f[0] for f in fw], dtype=np.float64), np.array([f[1] for f in fw], dtype=np.float64),
                         Px, Py, w, out=buf[:, a:b])
        a = b
    rays = RealRays(*[buf[q] for q in range(8)])
    sg = lens.surface_group
    sg.trace(rays, wavelengths=wls)
    lt = sg.last_trace
    if lt is not None and lt.out is not None and lt.out[0].data_ptr() == rays.x.data_ptr():
        out = lt.out
    else:
        out = torch.stack([rays.x, rays.y, rays.z, rays.L, rays.M, rays.N, rays.i, rays.opd])
    return rms_spot_sq(out, npair).sum()


def rms_merit_sharded(lens, fields, wavelengths, Px, Py, params=None):
    """BASELINE C4 across GPUs (SURVEY.md 8e level 1): the (field, wavelength) pairs of the merit are dealt round-robin
    to the ranks (`parallel.shard_items`), every rank traces ITS pairs forward + backward, and what crosses NVLink is
    one all-reduce of the loss scalar and -- with `params` -- ONE all-reduce (SUM) of the flat parameter-gradient
    vector after the local backward (the ranks hold different numbers of pairs, so per-trace collectives would not
    pair up; DDP's shape). Returns the GLOBAL loss (a detached scalar tensor); `params[i].grad` holds the global
    gradient on every rank, ready for the same optimiser step everywhere."""
    from . import parallel
    pairs = [(tuple(f), float(w)) for w in wavelengths for f in fields]
    mine = parallel.shard_items(pairs)
    dev = Px.device
    if mine:
        loss = _rms_merit_pairs(lens, mine, Px, Py)
        if params is not None:
            loss.backward()
        total = loss.detach().clone()
    else:
        total = torch.zeros((), dtype=torch.float64, device=dev)
    _, w = parallel.world()
    if w > 1:
        import torch.distributed as dist
        dist.all_reduce(total, op=dist.ReduceOp.SUM)
    if params is not None:
        parallel.allreduce_param_grads(params)
    return total


def smoke_backward() -> str:
    """forward + backward through the public API on the Cooke triplet: finite, non-zero gradients that agree
    with a central finite difference of the merit."""
    from . import backend as be
    from . import samples

    be.set_precision("float64")
    lens = samples.CookeTriplet()
    r1 = torch.tensor(22.01359, dtype=torch.float64, device=be.get_device(), requires_grad=True)
    lens.set_radius(r1, 1)
    rng = np.random.default_rng(0)
    rr, th = np.sqrt(rng.uniform(size=2000)), rng.uniform(0, 2 * np.pi, 2000)
    Px, Py = rr * np.cos(th), rr * np.sin(th)
    fields = [(0.0, 0.0), (0.0, 0.7), (0.0, 1.0)]
    loss = _rms_merit(lens, fields, 0.55, Px, Py)
    loss.backward()
    g = float(r1.grad)
    h = 1e-5
    vals = []
    for d in (+h, -h):
        lens.set_radius(22.01359 + d, 1)
        with torch.no_grad():
            vals.append(float(_rms_merit(lens, fields, 0.55, Px, Py)))
    fd = (vals[0] - vals[1]) / (2 * h)
    assert np.isfinite(g) and g != 0 and abs(g - fd) <= 1e-5 * abs(fd), (g, fd)
    return f"fwd+bwd ok: d(merit)/d(R1) = {g:.9e} (finite difference {fd:.9e})"


def bench_backward(lens, packed, ins64, timed, args, fp64_peak, hbm_peak=None) -> dict:
    """fwd + bwd on the bench batch (C2 lens, 50 M rays) and on the Cooke triplet (C1 / C4 lens, 9 x 1 M rays), kernel
    level: `trace_differentiable` = olrt_trace_fwd_f64 writing checkpoints + olrt_trace_bwd_f64, driven by a fixed
    d loss / d out. Legs: gradients w.r.t. the whole prescription AND the six ray components with [S][8][N] records
    (what a caller that also reads per-surface arrays gets); the same with tile-blocked checkpoints; and the BASELINE
    C4 configuration -- curvature / thickness / conic / asphere variables only: blocked checkpoints, no per-ray
    gradient arrays, no frame x / y partials (OLRT_BWD_NO_FRAME_XY). Forward + backward counts as 3x the forward's
    algorithmic flops."""
    from . import flops, samples
    from .prescription import pack_surface_group

    def no_decenter(pk_):   # what prescription_tensor records when no cs.x / cs.y is a variable
        pk_.live_offsets = frozenset(o for lay in pk_.layouts for o in range(lay.frames + 2, lay.aper))
        return pk_

    n = int(ins64[0].numel())
    S = packed.n_surfaces
    pk = pack_surface_group(lens.surface_group.surfaces, packed.wavelengths, record="all")
    pk.set_nr_iters(packed.nr_iters())
    dev = ins64[0].device
    presc = torch.tensor(pk.presc[0], dtype=torch.float64, device=dev, requires_grad=True)
    gout_w = torch.zeros((8, n), dtype=torch.float64, device=dev)
    gout_w[0] = 1.0
    gout_w[1] = 0.5
    fl = 3.0 * flops.flops_per_ray(pk)
    out = {}
    pkb = pack_surface_group(lens.surface_group.surfaces, packed.wavelengths, record="none")
    pkb.set_nr_iters(packed.nr_iters())
    pkc = no_decenter(pack_surface_group(lens.surface_group.surfaces, packed.wavelengths, record="none"))
    pkc.set_nr_iters(packed.nr_iters())
    n_ck = S
    legs = (("fp64_fwd_bwd", pk, True, False), ("fp64_fwd_bwd_blocked_checkpoints", pkb, True, True),
            ("fp64_fwd_bwd_prescription_only", pkc, False, True))
    for name, pk_, ray_grads, blocked in legs:
        rays = [t.detach().requires_grad_(ray_grads and q < 6) if t is not None else None
                for q, t in enumerate(ins64)]

        def step():
            presc.grad = None
            o, record, _ = trace_differentiable(pk_, presc, rays, fast=True, blocked=blocked)
            o.backward(gout_w)

        ms = timed(step, max(2, args.steps // 2), 3)
        # algorithmic HBM bytes per ray: fwd reads 7 inputs, writes 8 outputs + the checkpoints; bwd reads the
        # checkpoints, 7 inputs and 8 output gradients, writes 6 ray gradients when they are wanted
        by = 56 + 64 + 56 + 64 + (48 if ray_grads else 0) + 2 * (48 * (n_ck - 1) if blocked else 64 * S)
        out[name] = {"ray_surfaces_per_s": n * S / (ms * 1e-3), "ms": ms,
                     "frac_fp64_fma_peak": fl * n / (ms * 1e-3) / 1e12 / fp64_peak,
                     "hbm_bytes_per_ray": by, "hbm_gbs": by * n / (ms * 1e-3) / 1e9,
                     "frac_hbm_peak": (by * n / (ms * 1e-3) / 1e9 / hbm_peak) if hbm_peak else None,
                     "flops_per_ray_fwd_bwd": fl, "checkpoint_bytes_per_ray": 48 * n_ck if blocked else 64 * S,
                     "checkpoints": "tile-blocked [tile][surface][6][256] (x y z L M copysign(i, N))" if blocked else "[S][8][N] records",
                     "gradients": "prescription (frame z, radii, conics, asphere coefficients)" if not ray_grads else
                                  "prescription (frames, radii, conics, asphere coefficients) + x y z L M N of every ray",
                     "kernels": "trace_fx_kernel<double> writing the checkpoints + trace_bx_kernel<double> "
                                "(reduced-algebra adjoint)"}
    del gout_w, rays
    torch.cuda.empty_cache()
    # ---- Cooke triplet (north star: "the fused fp64 forward+backward trace on the Cooke triplet and ..."): 3 fields x
    # 3 wavelengths x 1 M rays in one batch with per-ray wavelengths, prescription gradients only
    cooke = samples.CookeTriplet()
    wls = [float(w) for w in cooke.wavelengths]
    fields = cooke.get_field_coords()
    n1 = 1_000_000
    rng = np.random.default_rng(0)
    r_, th = rng.uniform(size=n1), rng.uniform(0, 2 * np.pi, size=n1)
    Px = torch.from_numpy(np.sqrt(r_) * np.cos(th)).to(dev)
    Py = torch.from_numpy(np.sqrt(r_) * np.sin(th)).to(dev)
    nf, nw = len(fields), len(wls)
    Hx = torch.tensor([f[0] for f in fields], dtype=torch.float64, device=dev).repeat_interleave(n1).repeat(nw)
    Hy = torch.tensor([f[1] for f in fields], dtype=torch.float64, device=dev).repeat_interleave(n1).repeat(nw)
    wl = torch.tensor(wls, dtype=torch.float64, device=dev).repeat_interleave(nf * n1)
    crays = cooke.generate_rays(Hx, Hy, Px.repeat(nf * nw), Py.repeat(nf * nw), wl)
    cins = [t.detach() for t in crays.components()]
    nC = int(cins[0].numel())
    pkk = no_decenter(pack_surface_group(cooke.surface_group.surfaces, wls, record="none"))
    Sc = pkk.n_surfaces
    cpresc = torch.tensor(pkk.presc[0], dtype=torch.float64, device=dev, requires_grad=True)
    gw = torch.zeros((8, nC), dtype=torch.float64, device=dev)
    gw[0] = 1.0
    gw[1] = 0.5
    flc = 3.0 * flops.flops_per_ray(pkk)

    def cstep():
        cpresc.grad = None
        o, _, _ = trace_differentiable(pkk, cpresc, cins, fast=True, blocked=True)
        o.backward(gw)

    ms = timed(cstep, max(3, args.steps), 3)
    n_ckc = Sc
    out["fp64_fwd_bwd_cooke"] = {
        "ray_surfaces_per_s": nC * Sc / (ms * 1e-3), "ms": ms, "rays": nC,
        "frac_fp64_fma_peak": flc * nC / (ms * 1e-3) / 1e12 / fp64_peak, "flops_per_ray_fwd_bwd": flc,
        "checkpoint_bytes_per_ray": 48 * (n_ckc - 1),
        # 8 inputs (per-ray wavelengths) + 8 outputs + 6 of 7 checkpoint blocks written (the image plane has none),
        # then the checkpoints, 8 inputs and 8 output gradients read: the sphere-only triplet is HBM-bound
        "hbm_bytes_per_ray": 256 + 96 * (n_ckc - 1),
        "hbm_gbs": (256 + 96 * (n_ckc - 1)) * nC / (ms * 1e-3) / 1e9,
        "frac_hbm_peak": ((256 + 96 * (n_ckc - 1)) * nC / (ms * 1e-3) / 1e9 / hbm_peak) if hbm_peak else None,
        "bound": "hbm: 2586 flop over 832 B per ray = 3.1 flop/B, below the machine's 5.5 flop/B (35.8 TFLOP/s over "
                 "6.5 TB/s): the FP64 fraction is what the bandwidth leaves",
        "workload": "Cooke triplet, 3 fields x 3 wavelengths x 1 M random pupil rays (per-ray wavelengths), "
                    "prescription gradients (frame z, radii, conics), blocked checkpoints"}

    def cfwd():
        fused_trace(pkk, cins, fast=True, verify_nr=False, check_status=False, out=gw)

    ms = timed(cfwd, max(3, args.steps), 3)
    out["fp64_fwd_cooke"] = {"ray_surfaces_per_s": nC * Sc / (ms * 1e-3), "ms": ms, "rays": nC,
                             "frac_fp64_fma_peak": flc / 3.0 * nC / (ms * 1e-3) / 1e12 / fp64_peak,
                             "flops_per_ray": flc / 3.0,
                             "workload": "the same 9 M rays, forward only (ray arrays in, final state out)"}
    return out
