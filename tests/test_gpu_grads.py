"""Backward kernel against torch autograd on the reference's torch backend (tests/golden/*_grads.npz,
made by oracle/make_golden_grads.py). Bar: 1e-8 relative (BASELINE.json north_star), fp64."""
import numpy as np
import pytest
import torch

from helpers import GRAD_MIRRORS, load_golden, mirror_packed
from optiland_pr_b200.autograd import trace_differentiable

pytestmark = pytest.mark.gpu


def _run(name, dtype=torch.float64):
    g = load_golden(name + "_grads")
    lens, packed = mirror_packed(name, g["wavelengths"])
    assert np.array_equal(packed.ops, g["ops"]) and np.array_equal(packed.presc, g["presc"])
    presc = torch.tensor(g["presc"][0], dtype=dtype, device="cuda", requires_grad=True)
    rays = [torch.tensor(g["rays_in"][q], dtype=dtype, device="cuda", requires_grad=(q < 7)) for q in range(8)]
    out, record, _ = trace_differentiable(packed, presc, rays)
    w = torch.tensor(g["weights"], dtype=dtype, device="cuda")
    loss = (w * record).sum()
    loss.backward()
    gin = np.stack([rays[q].grad.cpu().numpy() for q in range(7)]).astype(np.float64)
    return g, float(loss), gin, presc.grad.cpu().numpy().astype(np.float64)


def _rel(a, b, floor):
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), floor)))


@pytest.mark.parametrize("name", sorted(GRAD_MIRRORS))
def test_fp64_gradients_match_reference_autograd(name):
    g, loss, gin, gpresc = _run(name)
    assert abs(loss - g["loss"]) <= 1e-10 * abs(g["loss"])
    for q in range(7):
        ref = g["gin"][q]
        assert _rel(gin[q], ref, 1e-3 * np.abs(ref).max() + 1e-300) <= 1e-8, f"d loss / d ray component {q}"
    have = g["have"]
    ref = g["gpresc"][have]
    got = gpresc[have]
    # every entry relative to itself; entries that are sums of cancelling per-ray terms get the floor of
    # 1e-6 x the largest gradient of the prescription
    assert _rel(got, ref, 1e-6 * np.abs(ref).max()) <= 1e-8, np.stack([got, ref], 1)
    # entries the reference has no leaf for stay zero
    assert np.all(gpresc[~have] == 0)
