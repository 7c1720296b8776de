"""GPU: batched prescriptions (Monte-Carlo tolerancing, BASELINE C5), the spot epilogue, and a torch optimiser
driving the fused forward + backward kernels (BASELINE C4)."""
import numpy as np
import pytest
import torch

from helpers import load_golden, mirror_packed
from oracle import trace_np
from optiland_pr_b200 import backend as be
from optiland_pr_b200 import samples, spot
from optiland_pr_b200.trace import fused_trace

pytestmark = pytest.mark.gpu


def test_monte_carlo_prescription_batch_matches_oracle_per_trial():
    """8 perturbed copies of the even-asphere double Gauss (radius 0.1 %, thickness 0.02 mm, decenter 0.1 mm)
    x 1024 rays each in ONE launch; every trial is compared with the oracle run on its own prescription."""
    g = load_golden("even_asphere_double_gauss")
    lens, packed = mirror_packed("even_asphere_double_gauss", g["wavelengths"], record="last")
    rng = np.random.default_rng(42)
    T, rpp = 8, 1024
    presc = np.repeat(packed.presc, T, axis=0)
    for t in range(1, T):
        for lay in packed.layouts:
            presc[t, lay.frames + 0] += rng.normal(0, 0.1)
            presc[t, lay.frames + 1] += rng.normal(0, 0.1)
            presc[t, lay.frames + 2] += rng.normal(0, 0.02)
            if np.isfinite(presc[t, lay.radius]):
                presc[t, lay.radius] *= 1.0 + rng.normal(0, 1e-3)
    base = g["rays_in"][:, :rpp]
    rays_in = np.tile(base, (1, T))
    rays = [torch.tensor(rays_in[q], device="cuda") for q in range(8)]
    # the batch must agree on the Newton trip counts: take the max over the trials' own counts (oracle)
    per_trial = []
    for t in range(T):
        packed_t = mirror_packed("even_asphere_double_gauss", g["wavelengths"], record="last")[1]
        packed_t.presc[0] = presc[t]
        per_trial.append(trace_np.trace_packed(packed_t, trace_np.Rays(*base)))
    counts = [max(o["nr_iters"][k] for o in per_trial) for k in range(packed.n_surfaces)]
    packed.set_nr_iters(counts)
    res = fused_trace(packed, rays, presc=torch.tensor(presc, device="cuda"), rays_per_presc=rpp, verify_nr=False)
    out = res.out.cpu().numpy()
    for t in range(T):
        packed_t = mirror_packed("even_asphere_double_gauss", g["wavelengths"], record="last")[1]
        packed_t.presc[0] = presc[t]
        ref = np.array(trace_np.trace_packed(packed_t, trace_np.Rays(*base), fixed_iters=counts)["records"][-1])
        got = out[:, t * rpp:(t + 1) * rpp]
        assert np.array_equal(np.isnan(got), np.isnan(ref))
        m = ~np.isnan(ref)
        assert np.max(np.abs(got[m] - ref[m]) / np.maximum(np.abs(ref[m]), 1.0)) <= 1e-10, t
    # trials differ from each other (the perturbations reached the kernel)
    assert not np.array_equal(out[:, :rpp], out[:, rpp:2 * rpp])


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
def test_spot_sums_against_numpy(dtype):
    rng = np.random.default_rng(3)
    G, n = 5, 40_000
    x = rng.normal(size=G * n) * 0.01 + np.repeat(np.arange(G), n) * 0.5
    y = rng.normal(size=G * n) * 0.02 - 1.0
    inten = (rng.uniform(size=G * n) > 0.1).astype(np.float64)
    xt, yt, it = (torch.tensor(a, dtype=dtype, device="cuda") for a in (x, y, inten))
    xs, ys = xt.double().cpu().numpy(), yt.double().cpu().numpy()
    tol = 1e-12 if dtype == torch.float64 else 1e-6
    # operand-style: all rays (ray.py:336-340)
    rms = spot.rms_spot_size(xt, yt, it, G).cpu().numpy()
    for g_ in range(G):
        sl = slice(g_ * n, (g_ + 1) * n)
        ref = np.sqrt(np.mean((xs[sl] - xs[sl].mean()) ** 2 + (ys[sl] - ys[sl].mean()) ** 2))
        assert abs(rms[g_] - ref) <= tol * max(ref, 1.0)
    # SpotDiagram-style: valid rays only, centroid, RMS and geometric radius (spot_diagram.py:317-357)
    r = spot.spot_radii(xt, yt, it, G)
    for g_ in range(G):
        sl = slice(g_ * n, (g_ + 1) * n)
        v = inten[sl] > 0
        cx, cy = xs[sl][v].mean(), ys[sl][v].mean()
        d2 = (xs[sl][v] - cx) ** 2 + (ys[sl][v] - cy) ** 2
        assert abs(float(r["cx"][g_]) - cx) <= tol and abs(float(r["cy"][g_]) - cy) <= tol
        assert abs(float(r["rms"][g_]) - np.sqrt(d2.mean())) <= 10 * tol
        assert abs(float(r["geo"][g_]) - np.sqrt(d2.max())) <= 10 * tol
        assert int(r["count"][g_]) == int(v.sum())


def test_nan_rays_poison_the_operand_like_the_reference():
    x = torch.tensor([0.0, 1.0, float("nan"), 2.0], dtype=torch.float64, device="cuda")
    rms = spot.rms_spot_size(x, x.clone(), torch.ones_like(x), 1)
    assert torch.isnan(rms).all()        # be.mean over an array with a NaN is NaN (ray.py:336-340)


def test_adam_reduces_rms_spot_merit_through_the_fused_kernels():
    """BASELINE C4 in miniature: radii of the Cooke triplet as torch Parameters, merit = sum of squared RMS
    spot sizes over the three fields, Adam steps through FusedTraceFn (forward + backward kernels)."""
    be.set_precision("float64")
    lens = samples.CookeTriplet()
    radii = {k: torch.nn.Parameter(torch.tensor(float(lens.surface_group.surfaces[k].geometry.radius) * 1.01,
                                                dtype=torch.float64, device="cuda")) for k in (1, 2, 5, 6)}
    for k, p in radii.items():
        lens.set_radius(p, k)
    rng = np.random.default_rng(0)
    rr, th = np.sqrt(rng.uniform(size=3000)), rng.uniform(0, 2 * np.pi, 3000)
    Px, Py = rr * np.cos(th), rr * np.sin(th)
    from optiland_pr_b200.autograd import _rms_merit
    fields = [(0.0, 0.0), (0.0, 0.7), (0.0, 1.0)]
    opt = torch.optim.Adam(list(radii.values()), lr=2e-2)
    history = []
    for _ in range(25):
        opt.zero_grad()
        loss = _rms_merit(lens, fields, 0.55, Px, Py)
        loss.backward()
        assert all(torch.isfinite(p.grad) and p.grad != 0 for p in radii.values())
        opt.step()
        history.append(float(loss))
    assert history[-1] < 0.5 * history[0], history


def test_fused_spot_merit_matches_the_torch_expression():
    """spot.rms_spot_sq (spot-sum kernel forward, analytic dense gradient backward) against the recorded torch
    expression of the same merit: value and d / d(radius, conic) through the public autograd path."""
    from optiland_pr_b200.autograd import _rms_merit, _rms_merit_fused
    be.set_precision("float64")
    rng = np.random.default_rng(2)
    rr, th = np.sqrt(rng.uniform(size=5000)), rng.uniform(0, 2 * np.pi, 5000)
    Px, Py = rr * np.cos(th), rr * np.sin(th)
    fields = [(0.0, 0.0), (0.0, 0.7), (0.0, 1.0)]
    got = []
    for merit in (_rms_merit, _rms_merit_fused):
        lens = samples.CookeTriplet()
        lens.surface_group.record = "none"
        r1 = torch.tensor(22.01359, dtype=torch.float64, device="cuda", requires_grad=True)
        k3 = torch.tensor(0.05, dtype=torch.float64, device="cuda", requires_grad=True)
        lens.set_radius(r1, 1)
        lens.set_conic(k3, 3)
        loss = merit(lens, fields, 0.55, Px, Py)
        loss.backward()
        got.append((float(loss.detach()), float(r1.grad), float(k3.grad)))
    (l0, a0, b0), (l1, a1, b1) = got
    assert abs(l1 - l0) <= 1e-10 * abs(l0), (l0, l1)
    assert abs(a1 - a0) <= 1e-8 * abs(a0) and abs(b1 - b0) <= 1e-8 * abs(b0), (got,)


def test_one_node_merit_over_all_wavelengths_matches_the_per_wavelength_sum():
    """_rms_merit_all (9 (wavelength, field) pairs, per-ray wavelengths, ONE FusedTraceFn node) against the sum of the
    per-wavelength merits: value and gradients."""
    from optiland_pr_b200.autograd import _rms_merit_all, _rms_merit_fused
    be.set_precision("float64")
    rng = np.random.default_rng(6)
    rr, th = np.sqrt(rng.uniform(size=2000)), rng.uniform(0, 2 * np.pi, 2000)
    Px = torch.from_numpy(rr * np.cos(th)).cuda()
    Py = torch.from_numpy(rr * np.sin(th)).cuda()
    got = []
    for one_node in (False, True):
        lens = samples.CookeTriplet()
        lens.surface_group.record = "none"
        fields, wls = lens.get_field_coords(), list(lens.wavelengths)
        r1 = torch.tensor(22.01359, dtype=torch.float64, device="cuda", requires_grad=True)
        r6 = torch.tensor(-18.39533, dtype=torch.float64, device="cuda", requires_grad=True)
        lens.set_radius(r1, 1)
        lens.set_radius(r6, 6)
        loss = (_rms_merit_all(lens, fields, wls, Px, Py) if one_node
                else sum(_rms_merit_fused(lens, fields, w, Px, Py) for w in wls))
        loss.backward()
        got.append((float(loss.detach()), float(r1.grad), float(r6.grad)))
    (l0, a0, b0), (l1, a1, b1) = got
    assert abs(l1 - l0) <= 1e-10 * abs(l0), got
    assert abs(a1 - a0) <= 1e-8 * abs(a0) and abs(b1 - b0) <= 1e-8 * abs(b0), got
