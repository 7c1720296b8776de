"""Fused ray launch + trace + spot epilogue (`olrt_trace_pupil_*`, include/olrt.h) against the separate kernels it
replaces: olrt_generate_rays_* (bit-exact vs the reference's RayGenerator, tests/test_raygen.py), olrt_trace_fwd_*
in tolerance mode (vs the reference: tests/test_gpu_golden.py, test_gpu_fuzz.py) and olrt_spot_sums_*."""
import numpy as np
import pytest
import torch

from oracle import raygen_np
from optiland_pr_b200 import _lib as L
from optiland_pr_b200 import backend as be
from optiland_pr_b200 import optic as optic_mod
from optiland_pr_b200 import samples, spot
from optiland_pr_b200.apertures import RadialAperture
from optiland_pr_b200.trace import fused_trace, fused_trace_pupil

pytestmark = pytest.mark.gpu

FIELDS = [(0.0, 0.0), (0.0, 0.7), (0.0, 1.0)]


def _setup(lens_cls, n_pupil, dtype, wild=False):
    be.set_precision("float64" if dtype == torch.float64 else "float32")
    lens = lens_cls()
    lens.surface_group.record = "none"
    if wild:  # a tight stop and an oversized pupil: vignetted rays, misses, TIR -> the IEEE fix-up pass
        lens.surface_group.surfaces[3].aperture = RadialAperture(r_max=4.0)
    wl = lens.primary_wavelength if hasattr(lens, "primary_wavelength") else lens.wavelengths[1]
    px, py = raygen_np.random_pupil(n_pupil, seed=3)
    if wild:
        px, py = px * 4.0, py * 4.0
    Px, Py = torch.tensor(px, device="cuda"), torch.tensor(py, device="cuda")
    Hx, Hy = np.array([f[0] for f in FIELDS]), np.array([f[1] for f in FIELDS])
    fxy, z_start = lens._field_offsets(Hx, Hy)
    packed = lens.surface_group.pack([float(wl)])
    kw = dict(EPD=lens._EPD, EPL=lens._EPL, z_start=z_start, wavelength=float(wl), dtype=dtype)
    return lens, packed, Px, Py, Hx, Hy, torch.tensor(fxy, device="cuda"), float(wl), kw


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("lens_cls,wild", [(samples.CookeTriplet, False), (samples.EvenAsphereDoubleGauss, False),
                                           (samples.EvenAsphereDoubleGauss, True)])
def test_pupil_mode_equals_launch_then_trace(lens_cls, wild, dtype):
    n_pupil = 70_001   # not a multiple of the CTA size: ragged last tile in every field
    lens, packed, Px, Py, Hx, Hy, fxy, wl, kw = _setup(lens_cls, n_pupil, dtype, wild)
    rays = lens.launch_rays(Hx, Hy, Px, Py, wl)
    ins = rays.components()
    ins[7] = None
    ref = fused_trace(packed, ins, fast=True, want_dir0=True)
    n = 3 * n_pupil
    out = torch.full((8, n), -7.0, dtype=dtype, device="cuda")
    dir0 = torch.full((3, n), -7.0, dtype=dtype, device="cuda")
    status = torch.zeros(1, dtype=torch.int32, device="cuda")
    # three ragged chunks cover the pupil
    for lo, hi in ((0, 1000), (1000, 33_333), (33_333, n_pupil)):
        fused_trace_pupil(packed, Px, Py, fxy, p_lo=lo, p_len=hi - lo, out=out, dir0=dir0, status_t=status, **kw)
    torch.cuda.synchronize()
    assert int(status.item()) == 0
    if wild:
        dead = torch.isnan(ref.out[0]).float().mean().item()
        assert 0.01 < dead < 0.99, dead
        assert (ref.out[6] == 0).any()
    assert torch.equal(torch.isnan(out), torch.isnan(ref.out))
    assert torch.equal(out.nan_to_num(0.0), ref.out.nan_to_num(0.0))
    assert torch.equal(dir0.nan_to_num(0.0), ref.dir0.nan_to_num(0.0))


@pytest.mark.parametrize("valid_only", [False, True])
@pytest.mark.parametrize("wild", [False, True])
def test_fused_spot_sums_equal_the_spot_kernel(wild, valid_only):
    n_pupil = 200_003
    lens, packed, Px, Py, Hx, Hy, fxy, wl, kw = _setup(samples.EvenAsphereDoubleGauss, n_pupil, torch.float64, wild)
    n = 3 * n_pupil
    out = torch.empty((8, n), dtype=torch.float64, device="cuda")
    shift = torch.tensor([[0.0, 0.0], [0.1, 12.0], [-0.2, 17.5]], dtype=torch.float64, device="cuda")
    sums = torch.zeros((3, L.SPOT_SUMS), dtype=torch.float64, device="cuda")
    sums_only = torch.zeros_like(sums)
    for lo, hi in ((0, 50_000), (50_000, n_pupil)):
        fused_trace_pupil(packed, Px, Py, fxy, p_lo=lo, p_len=hi - lo, out=out, spot=sums, spot_shift=shift,
                          spot_valid_only=valid_only, **kw)
        # epilogue only: nothing per ray is written
        fused_trace_pupil(packed, Px, Py, fxy, p_lo=lo, p_len=hi - lo, out=None, spot=sums_only, spot_shift=shift,
                          spot_valid_only=valid_only, **kw)
    want = spot.spot_sums(out[0], out[1], out[6], 3, n_pupil, shift=shift, valid_only=valid_only)
    torch.cuda.synchronize()
    got, got2, want = sums.cpu().numpy(), sums_only.cpu().numpy(), want.cpu().numpy()
    if wild and not valid_only:   # NaN rays poison the sums exactly like the reference's be.mean over all rays
        assert np.isnan(want[1:, 1]).any()
    for g_ in (got, got2):
        assert np.array_equal(g_[:, 0], want[:, 0])                       # counts
        assert np.array_equal(np.isnan(g_[:, :7]), np.isnan(want[:, :7]))
        m = ~np.isnan(want[:, :7])
        scale = np.maximum(np.abs(want[:, :7]), want[:, :1] * 1e-3)       # sums of ~n terms: relative to n * 1e-3
        assert np.all(np.abs(g_[:, :7][m] - want[:, :7][m]) <= 1e-11 * scale[m])
        assert np.allclose(g_[:, 5], want[:, 5], rtol=1e-14, atol=0, equal_nan=True)   # max r^2


def test_optic_trace_spot_matches_separate_epilogue():
    be.set_precision("float64")
    lens = samples.EvenAsphereDoubleGauss()
    lens.surface_group.record = "none"
    lens.pipeline_min_rays = 1 << 16
    px, py = raygen_np.random_pupil(262_144, seed=9)
    Px_h, Py_h = torch.from_numpy(px).pin_memory(), torch.from_numpy(py).pin_memory()
    Hx, Hy = [f[0] for f in FIELDS], [f[1] for f in FIELDS]
    rays = lens.trace(Hx, Hy, 0.5876, distribution=(Px_h, Py_h), spot=True)
    fused_rms = spot.spot_stats(rays.spot_sums)["rms"].cpu().numpy()
    sep_rms = spot.rms_spot_size(rays.x, rays.y, rays.i, 3, 262_144).cpu().numpy()
    assert np.allclose(fused_rms, sep_rms, rtol=1e-10, atol=0)
    x = rays.x.cpu().numpy().reshape(3, -1)
    y = rays.y.cpu().numpy().reshape(3, -1)
    ref = np.sqrt(((x - x.mean(1, keepdims=True)) ** 2 + (y - y.mean(1, keepdims=True)) ** 2).mean(1))
    assert np.allclose(fused_rms, ref, rtol=1e-10, atol=0)
    assert rays.L0 is not None and rays.w.shape == rays.x.shape


def test_unsupported_lens_is_refused_not_approximated():
    be.set_precision("float64")
    lens = samples.FreeformPlate("fringe")
    packed = lens.surface_group.pack([0.55])
    z = torch.zeros(4, dtype=torch.float64, device="cuda")
    fxy = torch.zeros((1, 2), dtype=torch.float64, device="cuda")
    out = torch.empty((8, 4), dtype=torch.float64, device="cuda")
    with pytest.raises(NotImplementedError):
        fused_trace_pupil(packed, z, z, fxy, out=out, EPD=1.0, EPL=0.0, z_start=-1.0, wavelength=0.55)


@pytest.mark.parametrize("npup,chunks", [(300_001, 5), (131_072, 12), (70_000, 3)])
def test_streamed_host_grid_equals_chunked_and_resident(npup, chunks):
    """Streamed mode (olrt_pupil.ready: ONE launch waits for the chunks the copy engine delivers) against the
    per-chunk launches and the device-resident grid: same bits, ragged last chunk and tile included."""
    be.set_precision("float64")
    px, py = raygen_np.random_pupil(npup, seed=21)
    Px_h, Py_h = torch.from_numpy(px).pin_memory(), torch.from_numpy(py).pin_memory()
    Hx, Hy = [f[0] for f in FIELDS], [f[1] for f in FIELDS]
    got = []
    for mode in ("streamed", "chunked", "resident"):
        lens = samples.EvenAsphereDoubleGauss()
        lens.surface_group.record = "none"
        lens.pipeline_min_rays = 1 << 12
        lens.pipeline_chunks = chunks
        lens.pipeline_streamed = (mode == "streamed")
        dist = (Px_h.cuda(), Py_h.cuda()) if mode == "resident" else (Px_h, Py_h)
        rays = lens.trace(Hx, Hy, 0.5876, distribution=dist, spot=True)
        got.append(rays)
        if mode != "resident":
            assert torch.equal(rays.spot_sums.cpu(), rays.spot_sums_host)
    for name in ("x", "y", "z", "L", "M", "N", "i", "opd", "L0", "M0", "N0"):
        assert torch.equal(getattr(got[0], name), getattr(got[1], name)), name
        assert torch.equal(getattr(got[0], name), getattr(got[2], name)), name
    sa, sb = spot.spot_stats(got[0].spot_sums), spot.spot_stats(got[2].spot_sums)
    for key in ("count", "cx", "cy", "rms", "intensity"):
        assert torch.allclose(sa[key], sb[key], rtol=1e-10, atol=1e-10), key


def test_streamed_launch_times_out_instead_of_hanging():
    """A "ready" word that is never set (a copy that was never queued) ends the launch with OLRT_ST_TIMEOUT after
    ~2 s; words that are set let the same launch through. The GPU is never left spinning."""
    import time
    _, packed, Px, Py, _, _, fxy, _, kw = _setup(samples.EvenAsphereDoubleGauss, 4096, torch.float64)
    out = torch.empty((8, 3 * 4096), dtype=torch.float64, device="cuda")
    for words, want in ((torch.ones(4, dtype=torch.int32, device="cuda"), 0),
                        (torch.tensor([1, 1, 0, 0], dtype=torch.int32, device="cuda"), L.ST_TIMEOUT)):
        status = torch.zeros(1, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        fused_trace_pupil(packed, Px, Py, fxy, out=out, status_t=status, ready=words, ready_chunk=1024, **kw)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        assert int(status.item()) & L.ST_TIMEOUT == want
        assert dt < 10.0
        if want:
            assert dt > 0.5


def test_optic_trace_raises_when_a_chunk_never_arrives():
    """The host side of the same failure: `Optic.trace` streams a pinned pupil grid; the words its copies write are
    made zero here (as if the copies had never landed), the kernel gives up after ~2 s with OLRT_ST_TIMEOUT and the
    call RAISES instead of returning rays computed from an unfilled buffer. The same lens then traces normally."""
    be.set_device("cuda:0")
    be.set_precision("float64")
    px, py = raygen_np.random_pupil(1 << 18, seed=5)      # >= 2 chunks of 65 536 points: the grid is streamed
    Px_h, Py_h = torch.from_numpy(px).pin_memory(), torch.from_numpy(py).pin_memory()
    lens = samples.EvenAsphereDoubleGauss()
    lens.surface_group.record = "none"
    lens.pipeline_min_rays = 1 << 12
    lens.pipeline_chunks = 4
    lens._ones_pinned = torch.zeros(64, dtype=torch.int32).pin_memory()
    with pytest.raises(L.OlrtLibraryError, match="OLRT_ST_TIMEOUT"):
        lens.trace([0.0], [1.0], 0.5876, distribution=(Px_h, Py_h))
    lens._ones_pinned = None
    rays = lens.trace([0.0], [1.0], 0.5876, distribution=(Px_h, Py_h))
    want = lens.trace([0.0], [1.0], 0.5876, distribution=(Px_h.cuda(), Py_h.cuda()))
    assert torch.equal(rays.x, want.x) and torch.equal(rays.opd, want.opd)


def test_multi_field_trace_uses_each_fields_own_newton_counts():
    """One Optic.trace over three fields == three single-field calls, bit for bit: the reference's Optic.trace takes one
    field per call, so every field's rays decide their own Newton exit test (newton_raphson.py:148). At this size the
    outer field needs one more update on one asphere than the inner two -- the one launch must not hand it to them."""
    be.set_device("cuda:0")
    be.set_precision("float64")
    px, py = optic_mod.random_pupil_device(1 << 22, 1, "cuda:0")
    Hx, Hy = [0.0, 0.0, 0.0], [0.0, 10.0 / 14.0, 1.0]
    lens = samples.EvenAsphereDoubleGauss()
    lens.surface_group.record = "none"
    got = lens.trace(Hx, Hy, 0.5876, distribution=(px, py), spot=True)
    n = px.numel()
    counts = []
    for f in range(3):
        one = samples.EvenAsphereDoubleGauss()
        one.surface_group.record = "none"
        want = one.trace([Hx[f]], [Hy[f]], 0.5876, distribution=(px, py), spot=True)
        counts.append(want.nr_iters_per_field[0])
        for name in ("x", "y", "z", "L", "M", "N", "i", "opd"):
            assert torch.equal(getattr(got, name)[f * n:(f + 1) * n], getattr(want, name)), (f, name)
        assert torch.allclose(got.spot_sums[f], want.spot_sums[0], rtol=1e-11, atol=1e-9)   # atomics: order of addition
    assert got.nr_iters_per_field == counts
    assert len({tuple(c) for c in counts}) > 1, "pick a size where the fields' counts differ"
    # the streamed host grid takes the same per-field counts
    lens2 = samples.EvenAsphereDoubleGauss()
    lens2.surface_group.record = "none"
    lens2.pipeline_chunks = 4
    s2 = lens2.trace(Hx, Hy, 0.5876, distribution=(px.cpu().pin_memory(), py.cpu().pin_memory()), spot=True)
    assert s2.nr_iters_per_field == counts
    assert torch.equal(s2.x, got.x) and torch.equal(s2.opd, got.opd)
