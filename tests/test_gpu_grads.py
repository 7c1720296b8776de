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
    _, packed = mirror_packed(name, g["wavelengths"])
    # a Zernike coefficient c_j sits in the row twice, as c_j (the slopes the reference codes, zernike.py:222-224) and as
    # c_j N_j (its sag, zernike/base.py:63-68): the kernel returns both partials, the coefficient's gradient is their chain
    gpresc = gpresc.copy()
    skip = np.zeros_like(have)
    for lay in packed.layouts:
        for (_, _, off, norm) in lay.zernike_terms:
            gpresc[off] += norm * gpresc[off + 1]
            skip[off + 1] = True
    ref = g["gpresc"][have]
    got = gpresc[have]
    assert _rel(got, ref, 1e-6 * np.abs(ref).max()) <= 1e-8, np.stack([got, ref], 1)
    trig = skip
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


def test_public_api_tilt_decenter_and_polynomial_variables_against_finite_differences():
    """`Optic.trace` under autograd on the three-mirror freeform system with live tensors on a tilt angle, a decenter
    and a polynomial coefficient (the mirror of optimization/variable/tilt.py, decenter.py, polynomial_coeff.py):
    d(RMS-spot merit) from the fused forward + general adjoint kernel against central finite differences of the same
    merit (an independent check of the whole chain: live tensors -> cos / sin rows -> kernel -> autograd)."""
    from optiland_pr_b200 import backend as be
    from optiland_pr_b200 import samples
    be.set_device("cuda:0")
    be.set_precision("float64")

    def merit(vals, grad):
        lens = samples.ThreeMirrorAnastigmat()
        sf = lens.surface_group.surfaces
        ps = [torch.tensor(v, dtype=torch.float64, device="cuda", requires_grad=grad) for v in vals]
        sf[2].geometry.cs.rx = ps[0]
        sf[3].geometry.cs.y = ps[1]
        co = [list(r) for r in sf[1].geometry.coefficients]
        co[2][1] = ps[2]
        sf[1].geometry.coefficients = co
        lens.surface_group.invalidate()
        total = 0.0
        with torch.set_grad_enabled(grad):
            for Hy in (0.0, 1.0, -1.0):
                rays = lens.trace([0.0], [Hy], 0.587, num_rays=6, distribution="hexapolar")
                total = total + ((rays.x - rays.x.mean()) ** 2 + (rays.y - rays.y.mean()) ** 2).mean()
        if grad:
            total.backward()
            return float(total), [float(p.grad) for p in ps]
        return float(total), None

    base = [-0.174532, -15.0, 2.97963986e-06]
    _, got = merit(base, True)
    for k, h in enumerate((1e-6, 1e-4, 1e-9)):
        up, dn = list(base), list(base)
        up[k] += h
        dn[k] -= h
        fd = (merit(up, False)[0] - merit(dn, False)[0]) / (2 * h)
        assert abs(got[k] - fd) <= 2e-6 * abs(fd), (k, got[k], fd)
