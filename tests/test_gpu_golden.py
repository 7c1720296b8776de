"""GPU parity against the reference's own outputs (tests/golden, made by oracle/make_golden.py).

Bar (BASELINE.json north_star): fp64 <= 1e-10 relative on positions and direction cosines with
bit-identical vignetting masks; fp32 <= 1e-5 relative.
"""
import numpy as np
import pytest
import torch

from helpers import CLOSED_FORM, MIRRORS, load_golden, mirror_packed, rel_err, fp32_errors
from optiland_pr_b200.trace import fused_trace

pytestmark = pytest.mark.gpu


def _trace(name, dtype, fast=False, replay=False):
    g = load_golden(name)
    lens, packed = mirror_packed(name, g["wavelengths"])
    rays = [torch.tensor(g["rays_in"][q], dtype=dtype, device="cuda") for q in range(8)]
    res = fused_trace(packed, rays, fast=fast, replay=replay)
    torch.cuda.synchronize()
    return g, res


@pytest.mark.parametrize("name", sorted(MIRRORS))
def test_fp64_strict_matches_reference(name):
    g, res = _trace(name, torch.float64)
    rec = res.record.cpu().numpy()
    ref = g["records"]
    assert rec.shape == ref.shape
    # vignetting masks bit-identical at every surface
    assert np.array_equal(rec[:, 6] == 0, ref[:, 6] == 0)
    for q in range(8):
        assert rel_err(rec[:, q], ref[:, q]) <= 1e-10, f"component {q}"
    if name in CLOSED_FORM:
        # + - * / sqrt only (intensity goes through exp): positions and directions bit for bit
        for q in (0, 1, 2, 3, 4, 5, 7):
            assert np.array_equal(rec[:, q], ref[:, q], equal_nan=True), f"component {q} not bit-exact"
    assert np.array_equal(res.out.cpu().numpy(), rec[-1], equal_nan=True)


@pytest.mark.parametrize("replay", [False, True], ids=["reduced_algebra", "replay_order"])
@pytest.mark.parametrize("name", sorted(MIRRORS))
def test_fp64_fast_within_tolerance(name, replay):
    """Tolerance mode (OLRT_FWD_FAST): <= 1e-10 relative on every component at every surface, vignetting / TIR /
    miss masks identical to the reference's. Lenses of planes, conics and even aspheres run the reduced-algebra
    kernel (csrc/trace_fx.cuh) unless `replay` asks for the reference's operation order."""
    g, res = _trace(name, torch.float64, fast=True, replay=replay)
    rec = res.record.cpu().numpy()
    ref = g["records"]
    assert np.array_equal(rec[:, 6] == 0, ref[:, 6] == 0)
    for q in range(8):
        assert rel_err(rec[:, q], ref[:, q]) <= 1e-10, f"component {q}"
    assert np.array_equal(res.out.cpu().numpy(), rec[-1], equal_nan=True)


@pytest.mark.parametrize("replay", [False, True], ids=["reduced_algebra", "replay_order"])
@pytest.mark.parametrize("name", sorted(MIRRORS))
def test_fp32_within_tolerance(name, replay):
    """fp32 bar: 1e-5 relative on position vectors and direction cosines (helpers.fp32_errors), masks
    identical. Per element with a 1 mm floor the bound asserted is 3e-5: a pure-fp32 evaluation of the
    reference's own formulas (NumPy float32) already sits at 1.3e-5 there, DESIGN.md "fp32 parity"."""
    g, res = _trace(name, torch.float32, replay=replay)
    rec = res.record.cpu().numpy().astype(np.float64)
    ref = g["records"]
    pos, dirs, opd = fp32_errors(rec, ref)
    assert pos <= 1e-5 and dirs <= 1e-5 and opd <= 1e-5, (pos, dirs, opd)
    for q in (0, 1, 2, 3, 4, 5, 7):
        assert rel_err(rec[:, q], ref[:, q]) <= 3e-5, f"component {q} (per element)"
    assert np.array_equal(rec[:, 6] == 0, ref[:, 6] == 0)
