"""CPU: the NumPy oracle reproduces the reference outputs stored in tests/golden (made by running the
unmodified reference, oracle/make_golden.py), and the mirror classes pack the same tables as the
reference's own objects."""
import numpy as np
import pytest

from helpers import FIELDS8, MIRRORS, load_golden, mirror_packed, rel_err
from oracle import trace_np


@pytest.mark.parametrize("name", sorted(MIRRORS))
def test_mirror_packs_like_reference_objects(name):
    g = load_golden(name)
    _, packed = mirror_packed(name, g["wavelengths"])
    assert np.array_equal(packed.ops, g["ops"])
    assert np.array_equal(packed.presc, g["presc"])


@pytest.mark.parametrize("name", sorted(MIRRORS))
def test_oracle_bit_exact_against_stored_reference(name):
    g = load_golden(name)
    _, packed = mirror_packed(name, g["wavelengths"])
    out = trace_np.trace_packed(packed, trace_np.Rays(*g["rays_in"]))
    rec = np.array(out["records"])
    assert rec.shape == g["records"].shape
    for q, f in enumerate(FIELDS8):
        assert np.array_equal(rec[:, q], g["records"][:, q], equal_nan=True), f
    if "nr_iters" in g:
        assert list(out["nr_iters"]) == list(g["nr_iters"])


@pytest.mark.parametrize("name", sorted(MIRRORS))
def test_fixed_iteration_replay_equals_exit_rule(name):
    """A launch with K_s fixed to the counts the exit rule produced gives the same bits
    (this is the contract between the kernel and the host driver)."""
    g = load_golden(name)
    _, packed = mirror_packed(name, g["wavelengths"])
    a = trace_np.trace_packed(packed, trace_np.Rays(*g["rays_in"]))
    b = trace_np.trace_packed(packed, trace_np.Rays(*g["rays_in"]), fixed_iters=a["nr_iters"])
    assert np.array_equal(np.array(a["records"]), np.array(b["records"]), equal_nan=True)
    assert rel_err(np.array(a["records"]), g["records"]) == 0.0


def test_wavefront_oracle_bit_exact():
    """oracle/wavefront_np.py against Wavefront(..., strategy="chief_ray").get_data of the unmodified reference
    (tests/golden/wavefront_cooke.npz, oracle/make_golden_wavefront.py): same bits for the OPD map in waves, the
    exit-pupil intersection points and the reference-sphere radius."""
    from oracle import wavefront_np
    g = load_golden("wavefront_cooke")
    names = ("x", "y", "z", "L", "M", "N", "opd")
    for k, f in enumerate(g["fields"]):
        rays = {n: g[f"f{k}_rays"][q] for q, n in enumerate(names)}
        rays["i"] = g[f"f{k}_rays"][7]
        chief = {n: g[f"f{k}_chief"][q] for q, n in enumerate(names)}
        got = wavefront_np.chief_ray_wavefront(rays, chief, tuple(f), float(g["wavelength"]), g["px"], g["py"],
                                               epd=float(g["EPD"]), max_field=float(g["max_field"]),
                                               pupil_z=float(g["pupil_z"]), n_image=float(g["n_image"]))
        assert got["radius"] == float(g[f"f{k}_radius"])
        assert np.array_equal(got["opd"], g[f"f{k}_opd"], equal_nan=True)
        assert np.array_equal(np.stack([got["pupil_x"], got["pupil_y"], got["pupil_z"]]), g[f"f{k}_pupil"], equal_nan=True)
        assert np.array_equal(got["intensity"], g[f"f{k}_intensity"])


@pytest.mark.parametrize("name", ["even_asphere_double_gauss", "three_mirror"])
def test_oracle_pool_reproduces_one_big_call(name):
    """oracle/pool.py (chunks on a process pool, batch-global Newton counts) == one call of the oracle on the batch."""
    from oracle import pool
    from helpers import mirror_packed
    g = load_golden(name)
    _, packed = mirror_packed(name, g["wavelengths"], record="none")
    whole = trace_np.Rays(*[a.copy() for a in g["rays_in"]])
    res = trace_np.trace_packed(packed, whole)
    got, k = pool.trace_batch(packed, g["rays_in"], chunk=97, procs=3)
    assert list(k) == list(res["nr_iters"])
    want = np.stack([whole.x, whole.y, whole.z, whole.L, whole.M, whole.N, whole.i, whole.opd])
    assert np.array_equal(got, want, equal_nan=True)
