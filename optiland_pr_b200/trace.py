"""Launch driver: torch tensors -> libolrt C ABI (include/olrt.h).

`fused_trace` is the one place that calls `olrt_trace_fwd_*`. It owns no numerics: every
arithmetic step of `Surface.trace` (optiland/surfaces/standard_surface.py:186-233) happens in
the CUDA kernel. PyTorch is used for device memory and streams only.

Newton-Raphson surfaces: the reference stops iterating when the maximum |f| over the WHOLE ray
batch drops below `tol` (optiland/geometries/newton_raphson.py:137-149), so every ray takes the
same number K_s of updates. The kernel takes K_s as an input and reports max|f| per iteration;
this driver checks the reported maxima against the reference's rule and relaunches with the
corrected K_s when the guess was wrong ("Newton contract").
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass

import numpy as np
import torch

from . import _lib as L
from .prescription import PackedSystem


@dataclass
class TraceResult:
    out: torch.Tensor            # [8, N]: x y z L M N intensity opd (global frame, after last surface)
    record: torch.Tensor | None  # [n_record, 8, N]
    dir0: torch.Tensor | None    # [3, N]: L0 M0 N0 at the last surface (local frame)
    nr_iters: list               # K_s actually used per intersected surface
    status: int
    launches: int
    fixups: int = -1             # rays re-traced by the IEEE fix-up kernel (fp64; -1 = not read back)
    nr_true: list | None = None  # with nr_floor: the counts this slice alone satisfies the exit rule with
    ckpt: torch.Tensor | None = None  # tile-blocked checkpoints (fused_trace(ckpt=True)), else None


def _row_ptrs(t: torch.Tensor):
    """Device addresses of the 8 rows of a contiguous [8, n] tensor (plain integer arithmetic: indexing the tensor
    eight times costs ~25 us of host time per launch)."""
    base, step = t.data_ptr(), t.stride(0) * t.element_size()
    return [base + q * step for q in range(8)]


def _ptr(t) -> int:
    return 0 if t is None else t.data_ptr()


_FIXUP = {}


def _fixup_workspace(n: int, device):
    """uint32 [1 + n] scratch for the IEEE re-trace list, cached per device (grown on demand)."""
    key = str(device)
    t = _FIXUP.get(key)
    if t is None or t.numel() < n + 1:
        t = torch.empty(max(n + 1, 1 << 16), dtype=torch.int32, device=device)
        _FIXUP[key] = t
    return t


_TABLE_CACHE: dict = {}      # (dtype, device, ops bytes, presc bytes) -> (ops, presc) device tensors, a few entries


def _device_tables(packed: PackedSystem, dtype, device):
    """ops / presc on the device. Cached by CONTENT (a few KB of bytes): a lens that is packed again with the same
    prescription -- every Optic.trace call packs -- does not upload its tables again."""
    key = (str(dtype), str(device), packed.ops.tobytes(), packed.presc.tobytes(), packed.presc.shape)
    hit = _TABLE_CACHE.get(key)
    if hit is not None:
        return hit
    ops = torch.from_numpy(packed.ops.copy()).to(device)
    presc = torch.from_numpy(np.ascontiguousarray(packed.presc)).to(device=device, dtype=dtype)
    if len(_TABLE_CACHE) >= 16:
        _TABLE_CACHE.pop(next(iter(_TABLE_CACHE)))
    _TABLE_CACHE[key] = (ops, presc)
    return ops, presc


def _true_iters(resid_row: np.ndarray, k_used: int, tol: float, max_iter: int):
    """Reference exit rule on the reported maxima. Returns (K_true or None if more are needed)."""
    for j in range(k_used + 1):
        v = resid_row[j]
        if math.isinf(v) or math.isnan(v):
            return max_iter  # a NaN ray never passes `max|f| < tol`: newton_raphson.py:148
        if j < max_iter and v < tol:
            return j
        if j == max_iter:
            return max_iter
    return None


def slice_true_iters(packed: PackedSystem, resid_rows: np.ndarray, k_used):
    """Reference exit rule on one slice's reported maxima (float64 view of a `resid_out` buffer).
    Returns the trip counts this slice satisfies the rule with, or None where it needs more than k_used."""
    out = list(k_used)
    for k in packed.nr_rows():
        lay = packed.layouts[k]
        out[k] = _true_iters(resid_rows[k], k_used[k], lay.tol_value, lay.max_iter)
    return out


def fused_trace(packed: PackedSystem, rays_in, opd_in=None, *, presc: torch.Tensor | None = None,
                n_presc: int = 1, rays_per_presc: int = 0, want_dir0: bool = False,
                fast: bool = False, verify_nr: bool = True, out: torch.Tensor | None = None,
                alias_out: bool = False, ckpt: bool = False, max_relaunch: int = 24, read_fixups: bool = True, check_status: bool = True,
                dir0_out: torch.Tensor | None = None, status_t: torch.Tensor | None = None,
                nr_floor: bool = False, resid_out: torch.Tensor | None = None,
                replay: bool = False, own_rows: bool | None = None) -> TraceResult:
    """Trace rays through a packed system on the current CUDA device and stream.

    rays_in: sequence of 8 tensors x y z L M N intensity wavelength (1-D, same length, CUDA,
             float64 or float32; wavelength may be None when the system has one wavelength).
    presc:   optional device tensor [n_presc, stride] overriding packed.presc (Monte Carlo batch).
    resid_out: zero-initialised int64 [n_surfaces, RESID_STRIDE] device tensor: the kernel reports its
             per-iteration residual maxima there and this call does NOT read them back (no host sync);
             the caller checks them later with `slice_true_iters` (slices of one batch, optic.py).
    fast:    tolerance mode (<= 1e-10 relative to the reference, identical masks) instead of bit-faithful;
             `replay` runs the tolerance build's straight-line / interpreter kernels instead of the reduced-algebra
             kernel even then (the reference's operation order for planes / conics / even aspheres, with FMA
             contraction; A/B measurements, fp32 too).
    own_rows: `presc` holds `packed`'s own row (same frame x / y entries), so the OLRT_HINT_ORDINARY judged on
             `packed` applies (default: only when `presc` is None).
    nr_floor: this call traces one SLICE of a larger batch: never lower the Newton trip counts below the
             ones in `packed` (another slice may need them), only raise them; the counts this slice alone
             would have needed come back in `TraceResult.nr_true`.
    """
    lib = L.load()
    x = rays_in[0]
    if not x.is_cuda:
        raise L.OlrtLibraryError("fused_trace needs CUDA tensors; there is no CPU path")
    dtype, device = x.dtype, x.device
    if dtype not in (torch.float64, torch.float32):
        raise TypeError(f"unsupported dtype {dtype}")
    n = int(x.numel())
    ins = []
    for q in range(8):
        t = rays_in[q] if q < len(rays_in) else None
        if t is None:
            if q < 7 or len(packed.wavelengths) > 1:
                raise ValueError(f"ray component {q} is required")
            ins.append(None)
            continue
        if t.dtype != dtype or t.device != device or t.numel() != n:
            raise ValueError("ray components must share dtype, device and length")
        ins.append(t.contiguous())
    if opd_in is not None:
        opd_in = opd_in.to(dtype).contiguous()
    ops_d, presc_d = _device_tables(packed, dtype, device)
    if presc is not None:
        presc_d = presc.to(device=device, dtype=dtype).contiguous()
        n_presc = int(presc_d.shape[0])
    n_rec = len(packed.record_surfaces)
    record = torch.empty((n_rec, 8, n), dtype=dtype, device=device) if n_rec else None
    ckpt_buf = None
    if ckpt:  # tile-blocked checkpoints [tile][surface][6][256] instead of records (include/olrt.h)
        if n_rec or not fast or n_presc != 1:
            raise L.OlrtLibraryError("blocked checkpoints need record='none', tolerance mode and one prescription")
        tiles = -(-n // L.CKPT_TILE)
        ckpt_buf = torch.empty(tiles * packed.n_surfaces * L.CKPT_VALS * L.CKPT_TILE, dtype=dtype, device=device)
    if out is None:
        if alias_out and n_rec and packed.record_surfaces[-1] == packed.layouts[-1].index:
            # the final state IS the last surface's record: both stores land on the same lines (merged in L2),
            # 64 B per ray less DRAM traffic and no second buffer
            out = record[n_rec - 1]
        else:
            out = torch.empty((8, n), dtype=dtype, device=device)
    dir0 = dir0_out if dir0_out is not None else (
        torch.empty((3, n), dtype=dtype, device=device) if want_dir0 else None)
    nr_rows = packed.nr_rows()
    own_status = status_t is None
    if own_status:  # a caller tracing one batch in slices shares one OR-accumulated word
        status_t = torch.zeros(1, dtype=torch.int32, device=device)
    nr_true = None
    fixup = _fixup_workspace(n, device) if dtype == torch.float64 else None
    fn = lib.olrt_trace_fwd_f64 if dtype == torch.float64 else lib.olrt_trace_fwd_f32
    in_arr = L.PtrArray8(*[_ptr(t) for t in ins])
    out_arr = L.PtrArray8(*_row_ptrs(out))
    # dir0 rows may be slices of a larger buffer: the kernel writes [3][n] with stride n, so a sliced
    # destination needs its own contiguous scratch (copied back below)
    dir0_tmp = None
    if dir0 is not None and not dir0.is_contiguous():
        dir0_tmp = torch.empty((3, n), dtype=dtype, device=device)
    dir0_ptr = _ptr(dir0_tmp if dir0_tmp is not None else dir0)
    stream = torch.cuda.current_stream(device).cuda_stream
    launches = 0
    with torch.cuda.device(device):
        while True:
            resid = None
            if nr_rows and resid_out is not None:
                resid = resid_out
            elif nr_rows and verify_nr:
                resid = torch.zeros((packed.n_surfaces, L.RESID_STRIDE), dtype=torch.int64, device=device)
                ops_d, _ = _device_tables(packed, dtype, device)
            sys_ = packed.make_sys(n_presc=n_presc, rays_per_presc=rays_per_presc,
                                   own_rows=(presc is None) if own_rows is None else own_rows)
            sys_.presc_stride = int(presc_d.shape[1])
            rc = fn(C.byref(sys_), presc_d.data_ptr(), ops_d.data_ptr(), n, in_arr, _ptr(opd_in),
                    out_arr, _ptr(ckpt_buf if ckpt else record), dir0_ptr, _ptr(resid), status_t.data_ptr(),
                    _ptr(fixup), (L.FWD_FAST if fast else 0) | (L.FWD_REPLAY if replay else 0)
                    | (L.FWD_CKPT_BLOCKED if ckpt else 0), stream)
            if rc == -4 and ckpt:
                raise NotImplementedError("olrt_trace_fwd: blocked checkpoints need a lens of planes / conics / even aspheres")
            L.check(rc, "olrt_trace_fwd")
            launches += 1
            if resid is None or n == 0 or resid_out is not None:
                break
            res = resid.cpu().numpy().view(np.float64)
            iters = packed.nr_iters()
            changed = False
            for k in nr_rows:
                lay = packed.layouts[k]
                kt = _true_iters(res[k], iters[k], lay.tol_value, lay.max_iter)
                if kt is None:
                    iters[k] = min(lay.max_iter, max(iters[k] * 2, iters[k] + 2))
                    changed = True
                    break
                if kt < iters[k] and nr_floor:
                    nr_true = nr_true if nr_true is not None else list(iters)
                    nr_true[k] = kt
                    continue
                if kt != iters[k]:
                    iters[k] = kt
                    changed = True
                    break
            if not changed:
                break
            if launches > max_relaunch:
                raise L.OlrtLibraryError("Newton iteration counts did not settle")
            packed.set_nr_iters(iters)
            nr_true = None
            if own_status:
                status_t.zero_()
    if dir0_tmp is not None:
        dir0.copy_(dir0_tmp)
    status = int(status_t.item()) if check_status else 0  # device->host sync
    if status & L.ST_BAD_HINT:
        raise L.OlrtLibraryError("kernel_hint does not match the ops table (packer bug)")
    fixups = int(fixup[0].item()) if (fixup is not None and read_fixups and check_status) else -1
    return TraceResult(out=out, record=record, dir0=dir0, nr_iters=packed.nr_iters(), status=status,
                       launches=launches, fixups=fixups, ckpt=ckpt_buf,
                       nr_true=nr_true if nr_true is not None else packed.nr_iters())


def fused_trace_pupil(packed: PackedSystem, Px: torch.Tensor, Py: torch.Tensor, field_xy: torch.Tensor, *,
                      EPD: float, EPL: float, z_start: float, wavelength: float, p_lo: int = 0,
                      p_len: int | None = None, dtype=torch.float64, out: torch.Tensor | None = None,
                      dir0: torch.Tensor | None = None, spot: torch.Tensor | None = None,
                      spot_shift: torch.Tensor | None = None, spot_valid_only: bool = False,
                      resid_out: torch.Tensor | None = None, status_t: torch.Tensor | None = None,
                      presc: torch.Tensor | None = None, group_k: torch.Tensor | None = None,
                      group_resid: torch.Tensor | None = None, field_vig: torch.Tensor | None = None,
                      field_z: torch.Tensor | None = None, ready: torch.Tensor | None = None,
                      ready_chunk: int = 0) -> None:
    """Enqueue `olrt_trace_pupil_*` (include/olrt.h): launch the rays of every field for pupil points
    [p_lo, p_lo + p_len) in the kernel, trace them, optionally accumulate the spot sums. Nothing is read back:
    the caller owns `status_t`, `resid_out` (see fused_trace) and checks them after the last chunk.
    out: [8, F * P] or None (epilogue only); dir0: [3, F * P] or None; spot: [F, SPOT_SUMS] float64 (zeroed by
    the caller); spot_shift: [F, 2] float64. presc: optional [n_presc, stride] device tensor (Monte-Carlo batch):
    every prescription traces the same fields and pupil range, F above becomes n_presc * F (trial-major).
    group_k: optional int32 [groups, n_surfaces] device tensor of Newton trip counts per group (trial x field)
    instead of the batch-global counts in `packed`; group_resid: int64 [groups, n_surfaces, GROUP_RESID] (zeroed)
    receives each group's residual maxima (`group_true_iters` applies the reference's exit rule to them).
    field_vig: optional [F, 2] float64 device tensor of (vx, vy) = 1 - vignetting factors; field_z: optional [F]
    float64 device tensor -- finite object, `field_xy` / `field_z` are then each field's object point (olrt.h).
    ready / ready_chunk: streamed pupil grid (include/olrt.h, olrt_pupil.ready): int32 device words, one per
    `ready_chunk` pupil points, that a copy engine sets as the host->device copy of Px / Py lands; the one launch
    covers the whole grid and waits, chunk by chunk, for the data.
    Raises NotImplementedError where the C ABI answers
    OLRT_E_UNSUPPORTED (the caller then uses launch + fused_trace + spot_sums)."""
    lib = L.load()
    if not Px.is_cuda:
        raise L.OlrtLibraryError("fused_trace_pupil needs CUDA tensors; there is no CPU path")
    device = Px.device
    npup = int(Px.numel())
    nf = int(field_xy.shape[0])
    p_len = npup - p_lo if p_len is None else p_len
    ops_d, presc_d = _device_tables(packed, dtype, device)
    n_presc = 1
    if presc is not None:
        presc_d = presc.to(device=device, dtype=dtype).contiguous()
        n_presc = int(presc_d.shape[0])
    pu = L.OlrtPupil(Px.data_ptr(), Py.data_ptr(), field_xy.data_ptr(), _ptr(field_vig), _ptr(field_z), npup, int(p_lo), int(p_len), nf, 0,
                     float(EPD), float(EPL), float(z_start), float(wavelength), _ptr(ready), int(ready_chunk))
    sys_ = packed.make_sys(n_presc=n_presc, rays_per_presc=nf * p_len, own_rows=presc is None)
    sys_.presc_stride = int(presc_d.shape[1])
    n_total = n_presc * nf * npup
    fixup = _fixup_workspace(n_total, device) if dtype == torch.float64 else None
    out_arr = None
    if out is not None:
        if out.dtype != dtype or tuple(out.shape) != (8, n_total) or not out.is_contiguous():
            raise ValueError("out must be a contiguous [8, n_presc * n_fields * n_pupil] tensor of the trace dtype")
        out_arr = C.byref(L.PtrArray8(*_row_ptrs(out)))
    fn = lib.olrt_trace_pupil_f64 if dtype == torch.float64 else lib.olrt_trace_pupil_f32
    with torch.cuda.device(device):
        rc = fn(C.byref(sys_), presc_d.data_ptr(), ops_d.data_ptr(), C.byref(pu), out_arr, _ptr(dir0),
                _ptr(resid_out), _ptr(status_t), _ptr(fixup), _ptr(spot), _ptr(spot_shift),
                L.SPOT_VALID_ONLY if spot_valid_only else 0, _ptr(group_k), _ptr(group_resid), L.FWD_FAST,
                torch.cuda.current_stream(device).cuda_stream)
    if rc == -4:
        raise NotImplementedError("olrt_trace_pupil: this lens / mode needs the separate launch + trace kernels")
    L.check(rc, "olrt_trace_pupil")


def group_true_iters(packed: PackedSystem, resid: np.ndarray, k_used: np.ndarray) -> np.ndarray:
    """The reference's exit rule (newton_raphson.py:137-149) per group. resid: float64 view [G, S, GROUP_RESID] of
    a `group_resid` buffer; k_used: int [G, S]. Returns int [G, S]: the trip counts each group satisfies the rule
    with, -1 where it needs more than it was given (or more than GROUP_RESID - 1)."""
    G = resid.shape[0]
    out = np.zeros((G, packed.n_surfaces), dtype=np.int32)
    for k in packed.nr_rows():
        lay = packed.layouts[k]
        r = resid[:, k, :]
        bad = ~np.isfinite(r)
        ok = (r < lay.tol_value) & ~bad
        j = np.arange(L.GROUP_RESID)[None, :]
        seen = j <= k_used[:, k][:, None]                 # slots the kernel actually reported
        hit = ok & seen & (j < lay.max_iter)
        first = np.where(hit.any(axis=1), hit.argmax(axis=1), -1)
        first = np.where((bad & seen).any(axis=1), -1, first)   # a NaN ray: the reference runs max_iter
        out[:, k] = first
    return out


def fused_spot_groups(packed: PackedSystem, presc_t: torch.Tensor, Px: torch.Tensor, Py: torch.Tensor,
                      field_xy: torch.Tensor, gen_kw: dict, k0: np.ndarray, *, shift: torch.Tensor | None = None,
                      valid_only: bool = False, settle: bool = True, between=None, keep: torch.Tensor | None = None):
    """Spot sums of every (prescription, field) group from fused launch + trace + spot kernels that write nothing
    per ray. `presc_t` [T, stride] holds one prescription per row (Monte-Carlo trials, or one row per WAVELENGTH:
    the media live in the prescription row); the groups are g = row * n_fields + field. The Newton trip counts are
    PER GROUP, as in the reference where every (trial / wavelength, field) is its own trace (newton_raphson.py:148):
    `k0` [n_fields, S] is the first guess, groups whose counts fail the exit rule are traced again with corrected
    counts (settle=False: `k0` is [G, S] and trusted). shift: [G, 2] point the sums are taken about (default: each
    group's chief-ray image point, from one more tiny launch). Returns (sums [G, SPOT_SUMS] with columns 8:10 =
    shift, counts [G, S] int32). `between`: optional host callback run ONCE right after the first launch is queued and
    before its results are waited for (the caller's host work then overlaps the kernel). `keep`: optional [8, G * P]
    buffer that receives the final ray states of the FIRST launch (all groups); `keep.valid` is set True when that
    launch's counts passed the exit rule for every group (so the buffer holds the result), else False."""
    dev = Px.device
    dt = gen_kw.get("dtype", torch.float64)
    nf, S = int(field_xy.shape[0]), packed.n_surfaces
    T = int(presc_t.shape[0])
    G = T * nf
    nr_rows = packed.nr_rows()
    kg = (np.repeat(k0[None, :, :], T, axis=0).reshape(G, S) if k0.shape[0] == nf and k0.ndim == 2 and G != nf
          else np.asarray(k0).reshape(G, S)).astype(np.int32)
    packed.set_nr_iters(np.maximum(kg.max(axis=0), 0).tolist())   # (see the launch below)
    if shift is None:
        zero1 = torch.zeros(1, dtype=torch.float64, device=dev)
        chief = torch.empty((8, G), dtype=dt, device=dev)
        fused_trace_pupil(packed, zero1, zero1, field_xy, out=chief, presc=presc_t,
                          group_k=torch.from_numpy(kg).to(dev), **gen_kw)
        shift = chief[:2].t().to(torch.float64).nan_to_num(0.0, 0.0, 0.0).contiguous()
    out_sums = torch.empty((G, L.SPOT_SUMS), dtype=torch.float64, device=dev)
    todo = np.arange(T)
    keep_ok = keep is not None
    for round_ in range(8):
        sel = torch.from_numpy(todo).to(dev)
        gsel = (todo[:, None] * nf + np.arange(nf)[None, :]).reshape(-1)
        gsel_t = torch.from_numpy(gsel).to(dev)
        k_t = torch.from_numpy(kg[gsel]).to(dev)
        sums = torch.zeros((gsel.size, L.SPOT_SUMS), dtype=torch.float64, device=dev)
        check = bool(nr_rows) and settle
        gres = torch.zeros((gsel.size, S, L.GROUP_RESID), dtype=torch.int64, device=dev) if check else None
        sh = shift[gsel_t].contiguous()
        # the kernel lays its residual slots out by the ops table's counts and never gives a group more than those
        packed.set_nr_iters(np.maximum(kg[gsel].max(axis=0), 0).tolist())
        fused_trace_pupil(packed, Px, Py, field_xy, out=keep if (keep is not None and round_ == 0) else None, spot=sums,
                          spot_shift=sh, spot_valid_only=valid_only,
                          presc=presc_t[sel].contiguous(), group_k=k_t, group_resid=gres, **gen_kw)
        if round_ > 0:
            keep_ok = False       # some group was traced again with other counts: the kept states are stale
        sums[:, 8:10] = sh
        out_sums[gsel_t] = sums
        if between is not None:
            between()
            between = None
        if not check:
            break
        kt = group_true_iters(packed, gres.cpu().numpy().view(np.float64), kg[gsel])
        wrong = (kt != kg[gsel]).any(axis=1)
        if not wrong.any():
            break
        need_more = kt < 0
        if (need_more & (kg[gsel] >= L.GROUP_RESID - 1)).any():
            raise NotImplementedError("a group needs more than %d Newton updates" % (L.GROUP_RESID - 1))
        kg[gsel] = np.where(need_more, np.minimum(kg[gsel] * 2 + 1, L.GROUP_RESID - 1), kt)
        todo = np.unique(gsel[wrong] // nf)
    else:
        raise L.OlrtLibraryError("Newton trip counts did not settle")
    if keep is not None:
        keep.valid = bool(keep_ok)
    return out_sums, kg
