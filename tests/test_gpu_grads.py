"""Backward kernel against torch autograd on the reference's torch backend (tests/golden/*_grads.npz,
made by oracle/make_golden_grads.py). Bar: 1e-8 relative (BASELINE.json north_star), fp64."""
import numpy as np
import pytest
import torch

from helpers import GRAD_MIRRORS, GRAD_MIRRORS_GENERAL, load_golden, mirror_packed
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


@pytest.mark.parametrize("name", sorted(GRAD_MIRRORS_GENERAL))
def test_fp64_gradients_of_tilted_and_polynomial_surfaces_match_reference_autograd(name):
    """SURVEY.md 8f / VERDICT r1 item 7: the general adjoint kernel on tilted, decentred frames and 4 x 4 polynomial
    freeform mirrors (BASELINE C3's three-mirror system). Same bar: 1e-8 against torch autograd on the reference.
    d loss / d angle is chained from the kernel's d loss / d cos, d sin entries of the frame."""
    g, loss, gin, gpresc = _run(name)
    assert abs(loss - g["loss"]) <= 1e-10 * abs(g["loss"])
    for q in range(7):
        ref = g["gin"][q]
        assert _rel(gin[q], ref, 1e-3 * np.abs(ref).max() + 1e-300) <= 1e-8, f"d loss / d ray component {q}"
    have = g["have"]
    ref = g["gpresc"][have]
    got = gpresc[have]
    assert _rel(got, ref, 1e-6 * np.abs(ref).max()) <= 1e-8, np.stack([got, ref], 1)
    _, packed = mirror_packed(name, g["wavelengths"])
    trig = np.zeros_like(have)
    n_checked = 0
    for k, lay in enumerate(packed.layouts):
        for q in range(3):                                   # rz, ry, rx: cos at frames + 3 + 2q, sin at + 4 + 2q
            oc, os_ = lay.frames + 3 + 2 * q, lay.frames + 4 + 2 * q
            trig[oc] = trig[os_] = True
            if g["have_angle"][k, q]:
                c, s_ = g["presc"][0][oc], g["presc"][0][os_]
                got_a = -s_ * gpresc[oc] + c * gpresc[os_]
                ref_a = g["gangle"][k, q]
                assert abs(got_a - ref_a) <= 1e-8 * max(abs(ref_a), 1e-6 * np.abs(g["gangle"]).max()), (k, q, got_a, ref_a)
                n_checked += 1
    assert n_checked == int(g["have_angle"].sum()) > 0
    assert np.all(gpresc[~have & ~trig] == 0)
