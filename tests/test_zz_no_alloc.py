"""olrt_trace_fwd_noalloc_f64 (include/olrt.h): the forward trace whose re-trace list is a static device array
inside libolrt.so, for callers that cannot allocate. Called through the C ABI (ctypes) and held against
olrt_trace_fwd_f64 on the same inputs: the kernels are the same, so every output must be the same bits.
Two calls in a row share the static array; the first leaves entries in it (dead rays), the second must not see them.
Named test_zz_* so pytest collects it last."""
import ctypes as C

import numpy as np
import pytest
import torch

from helpers import load_golden, mirror_packed
from optiland_pr_b200 import _lib as L
from optiland_pr_b200.trace import _device_tables, fused_trace

pytestmark = pytest.mark.gpu


def _call(name, packed, rays_np, fast, with_fixup):
    """One trace through the C ABI; returns (out [8, n], record [S, 8, n], status, re-traced rays or None)."""
    lib = L.load()
    dev = torch.device("cuda", torch.cuda.current_device())
    n = rays_np.shape[1]
    rays = [torch.tensor(rays_np[q], dtype=torch.float64, device=dev) for q in range(8)]
    ops_d, presc_d = _device_tables(packed, torch.float64, dev)
    n_rec = len(packed.record_surfaces)
    out = torch.full((8, n), -7.0, dtype=torch.float64, device=dev)
    record = torch.full((n_rec, 8, n), -7.0, dtype=torch.float64, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    nr_rows = packed.nr_rows()
    resid = torch.zeros((packed.n_surfaces, L.RESID_STRIDE), dtype=torch.int64, device=dev) if nr_rows else None
    sys_ = packed.make_sys(n_presc=1, rays_per_presc=0)
    in_arr = L.PtrArray8(*[t.data_ptr() for t in rays])
    out_arr = L.PtrArray8(*[out[q].data_ptr() for q in range(8)])
    stream = torch.cuda.current_stream(dev).cuda_stream
    flags = L.FWD_FAST if fast else 0
    common = (C.byref(sys_), presc_d.data_ptr(), ops_d.data_ptr(), n, in_arr, None, out_arr, record.data_ptr(), None,
              resid.data_ptr() if resid is not None else None, status.data_ptr())
    fixup = None
    if with_fixup:
        fixup = torch.full((n + 1,), 0x55555555, dtype=torch.int32, device=dev)
        rc = getattr(lib, name)(*common, fixup.data_ptr(), flags, stream)
    else:
        rc = getattr(lib, name)(*common, flags, stream)
    assert rc == 0, (name, L.ERRORS.get(rc, rc))
    torch.cuda.synchronize()
    return (out.cpu().numpy(), record.cpu().numpy(), int(status.item()),
            int(fixup[0].item()) if fixup is not None else None)


def _same(a, b):
    return a.shape == b.shape and a.tobytes() == b.tobytes()   # bit for bit, NaN payloads included


CASES = [
    ("freeform_plate_fringe", False),   # interpreter kernel, TIR on part of the pupil: the list is NOT empty
    ("cooke_triplet", True),            # straight-line kernel, tolerance mode, no dead rays
    ("even_asphere_double_gauss", True),
    ("freeform_plate_fringe", True),
    ("cooke_triplet", False),           # strict kernel
]


def test_noalloc_matches_the_allocating_entry_point_on_calls_in_a_row():
    dead_total = 0
    for name, fast in CASES:   # consecutive calls share the static array
        g = load_golden(name)
        _, packed = mirror_packed(name, g["wavelengths"])
        rays_np = np.ascontiguousarray(g["rays_in"][:8], dtype=np.float64)
        # settle the Newton trip counts in `packed` (the C ABI takes them as input)
        fused_trace(packed, [torch.tensor(rays_np[q], device="cuda") for q in range(8)], fast=fast)
        want_out, want_rec, want_st, n_fix = _call("olrt_trace_fwd_f64", packed, rays_np, fast, True)
        for rep in range(2):
            got_out, got_rec, got_st, _ = _call("olrt_trace_fwd_noalloc_f64", packed, rays_np, fast, False)
            assert got_st == want_st, (name, fast, rep)
            assert _same(got_out, want_out), (name, fast, rep)
            assert _same(got_rec, want_rec), (name, fast, rep)
            assert not (got_out == -7.0).all(axis=0).any(), "a ray was left unwritten"
        # the allocating entry point agrees with the stored outputs of the reference (so both do)
        m = ~np.isnan(g["records"])
        assert np.array_equal(np.isnan(want_rec), ~m)
        assert np.max(np.abs(want_rec[m] - g["records"][m]) / np.maximum(np.abs(g["records"][m]), 1.0)) <= 1e-10
        dead_total += n_fix
    assert dead_total > 0, "no case exercised the re-trace list"


def test_noalloc_after_a_longer_list_and_ragged_sizes():
    """A call with many dead rays fills the list; shorter calls after it must not re-trace stale entries."""
    g = load_golden("freeform_plate_fringe")
    _, packed = mirror_packed("freeform_plate_fringe", g["wavelengths"])
    base = np.ascontiguousarray(g["rays_in"][:8], dtype=np.float64)
    reps = 40
    big = np.tile(base, (1, reps))
    want = _call("olrt_trace_fwd_f64", packed, big, True, True)
    got = _call("olrt_trace_fwd_noalloc_f64", packed, big, True, False)
    assert want[3] > 0 and _same(got[0], want[0]) and _same(got[1], want[1])
    for n in (1, 31, 257, base.shape[1]):
        small = np.ascontiguousarray(base[:, :n])
        want = _call("olrt_trace_fwd_f64", packed, small, True, True)
        got = _call("olrt_trace_fwd_noalloc_f64", packed, small, True, False)
        assert _same(got[0], want[0]) and _same(got[1], want[1]), n


def test_noalloc_rejects_sizes_beyond_its_static_array():
    lib = L.load()
    g = load_golden("cooke_triplet")
    _, packed = mirror_packed("cooke_triplet", g["wavelengths"])
    sys_ = packed.make_sys(n_presc=1, rays_per_presc=0)
    arr = L.PtrArray8()
    rc = lib.olrt_trace_fwd_noalloc_f64(C.byref(sys_), 1, 1, L.NOALLOC_MAX_RAYS + 1, arr, None, arr, None, None,
                                        None, None, L.FWD_FAST, None)
    assert rc == -2   # OLRT_E_LIMIT, before any CUDA call
    # an empty batch is accepted like the allocating entry point accepts it
    assert lib.olrt_trace_fwd_noalloc_f64(C.byref(sys_), 1, 1, 0, arr, None, arr, None, None, None, None,
                                          L.FWD_FAST, None) == 0
