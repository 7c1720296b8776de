"""TEST INFRASTRUCTURE ONLY -- gradient fixtures from torch autograd on the UNMODIFIED reference.

    python oracle/make_golden_grads.py        # needs /root/reference (build container only)

For each lens: the reference is switched to its torch backend (cpu, float64, grad enabled), every
differentiable prescription entry of the hot path (geometry.radius, geometry.k, even-asphere and
polynomial coefficients, cs.x / cs.y / cs.z and the non-zero tilt angles cs.rx / ry / rz of every
surface) and every input-ray component is made a leaf
tensor, `surface_group.trace(rays)` runs, and the scalar
        loss = sum_s sum_q sum_rays  w[s, q, ray] * record[s, q, ray]       (w: fixed random weights)
is back-propagated. Stored: rays_in, w, the loss, d loss / d rays_in ([7, N]: x y z L M N i) and
d loss / d presc in the packed layout of optiland_pr_b200/prescription.py (zeros where the reference
has no differentiable leaf: media, tol) and `gangle` [S, 3] = d loss / d (rz, ry, rx) of every surface's
frame -- the kernel returns d loss / d cos and d sin of each angle, the test chains them. The -m gpu
tests feed the same rays and w to the CUDA backward kernel and compare at 1e-8 (BASELINE.json north_star).
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_shim  # noqa: E402

ref_shim.install()

import torch  # noqa: E402

import optiland.backend as be  # noqa: E402
from optiland.rays import RealRays  # noqa: E402
from optiland.samples.objectives import CookeTriplet, DoubleGauss  # noqa: E402
from optiland.samples.simple import AsphericSinglet  # noqa: E402

from oracle.make_golden import ref_even_asphere_double_gauss  # noqa: E402
from oracle import raygen_np  # noqa: E402
from optiland_pr_b200 import _lib as L  # noqa: E402
from optiland_pr_b200.prescription import pack_surface_group  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")
FIELDS8 = ("x", "y", "z", "L", "M", "N", "intensity", "opd")


def leafify(lens):
    """Replace every differentiable hot-path parameter by a float64 leaf tensor. Returns
    [(surface index k (intersected numbering), name, sub-index, tensor)]."""
    leaves = []
    mk = lambda v: torch.tensor(float(v), dtype=torch.float64, requires_grad=True)
    for k, surf in enumerate(lens.surface_group.surfaces[1:]):
        g = surf.geometry
        cs = g.cs
        for nm in ("x", "y", "z"):
            t = mk(getattr(cs, nm))
            setattr(cs, nm, t)
            leaves.append((k, "frame_" + nm, 0, t))
        for q, nm in enumerate(("rz", "ry", "rx")):      # tilts (optimization/variable/tilt.py:59): truthy angles only,
            if float(getattr(cs, nm)) != 0.0:            # the reference skips a rotation whose angle is falsy
                t = mk(getattr(cs, nm))
                setattr(cs, nm, t)
                leaves.append((k, "angle", q, t))
        if type(g).__name__ == "Plane":
            continue
        if np.isfinite(float(g.radius)):
            g.radius = mk(g.radius)
            leaves.append((k, "radius", 0, g.radius))
        g.k = mk(g.k)
        leaves.append((k, "conic", 0, g.k))
        if hasattr(g, "coefficients") and type(g).__name__ in ("EvenAsphere", "OddAsphere"):
            cs_ = [mk(c) for c in g.coefficients]
            g.coefficients = cs_
            for i, c in enumerate(cs_):
                leaves.append((k, "coef", i, c))
        if type(g).__name__ == "PolynomialGeometry":     # C[i][j], optimization/variable/polynomial_coeff.py:72
            c2 = torch.tensor(np.asarray(be.to_numpy(g.coefficients), dtype=np.float64), requires_grad=True)
            g.coefficients = c2
            leaves.append((k, "coef2d", 0, c2))
        if type(g).__name__ == "ZernikePolynomialGeometry":   # optimization/variable/zernike_coeff.py:71
            cz = torch.tensor(np.asarray(be.to_numpy(g.zernike.coeffs), dtype=np.float64), requires_grad=True)
            g.zernike.coeffs = cz
            leaves.append((k, "zernike", 0, cz))
    return leaves


def run_case(name, make_lens, wavelengths, n_pupil, seed, pupil_scale=0.9, grad_mode=True):
    # rays and packed layout from the NumPy backend (identical inputs for both backends)
    be.set_backend("numpy")
    lens = make_lens()
    fields = lens.fields.get_field_coords()
    rng = np.random.default_rng(seed)
    parts = []
    for (hx, hy) in fields:
        for w in wavelengths:
            px, py = raygen_np.random_pupil(n_pupil, int(rng.integers(1 << 30)))
            r = lens.ray_tracer.ray_generator.generate_rays(np.full(n_pupil, hx), np.full(n_pupil, hy), px * pupil_scale, py * pupil_scale, w)
            parts.append(np.stack([np.asarray(getattr(r, q), dtype=np.float64) for q in ("x", "y", "z", "L", "M", "N", "i", "w")]))
    rays_in = np.concatenate(parts, axis=1)
    packed = pack_surface_group(lens.surface_group.surfaces, wavelengths)
    n = rays_in.shape[1]
    S = packed.n_surfaces
    w8 = rng.normal(size=(S, 8, n))

    be.set_backend("torch")
    be.set_device("cpu")
    be.set_precision("float64")
    # (grad mode makes every be.array a requires-grad leaf; the reference's Zernike radial term then floor-divides such
    # tensors -- (n + m) // 2, zernike/base.py -- which torch cannot differentiate: those lenses run with grad mode off
    # and only the explicit leaves below asking for gradients)
    if grad_mode:
        be.grad_mode.enable()
    else:
        be.grad_mode.disable()
    lens_t = make_lens()
    leaves = leafify(lens_t)
    comps = [torch.tensor(rays_in[q], dtype=torch.float64, requires_grad=(q < 7)) for q in range(8)]
    rays = RealRays(*comps)
    sg = lens_t.surface_group
    sg.trace(rays)
    wt = torch.tensor(w8)
    loss = 0
    for s, surf in enumerate(sg.surfaces[1:]):
        for q, f in enumerate(FIELDS8):
            loss = loss + (wt[s, q] * getattr(surf, f)).sum()
    loss.backward()
    gin = np.stack([comps[q].grad.numpy() if comps[q].grad is not None else np.zeros(n) for q in range(7)])
    gpresc = np.zeros(packed.stride)
    have = np.zeros(packed.stride, dtype=bool)
    gangle = np.zeros((S, 3))
    have_angle = np.zeros((S, 3), dtype=bool)
    for (k, nm, i, t) in leaves:
        lay = packed.layouts[k]
        if nm == "angle":
            gangle[k, i] = 0.0 if t.grad is None else float(t.grad)
            have_angle[k, i] = True
            continue
        if nm == "zernike":     # d loss / d c_j stored at the term's `c` entry (the kernel splits it over c and c N)
            gz = np.zeros(tuple(t.shape)) if t.grad is None else t.grad.numpy()
            for j, (_, _, off, _) in enumerate(lay.zernike_terms):
                gpresc[off] = gz[j]
                have[off] = True
            continue
        if nm == "coef2d":
            g2 = np.zeros(tuple(t.shape)) if t.grad is None else t.grad.numpy()
            gpresc[lay.coef:lay.coef + g2.size] = g2.reshape(-1)
            have[lay.coef:lay.coef + g2.size] = True
            continue
        off = {"frame_x": lay.frames, "frame_y": lay.frames + 1, "frame_z": lay.frames + 2,
               "radius": lay.radius, "conic": lay.conic, "coef": lay.coef + i}[nm]
        gpresc[off] = 0.0 if t.grad is None else float(t.grad)
        have[off] = True
    be.grad_mode.disable()
    be.set_backend("numpy")
    out = dict(rays_in=rays_in, weights=w8, loss=float(loss.detach()), gin=gin, gpresc=gpresc, have=have,
               gangle=gangle, have_angle=have_angle,
               wavelengths=np.asarray(wavelengths, dtype=np.float64), ops=packed.ops, presc=packed.presc)
    path = os.path.join(GOLDEN, name + "_grads.npz")
    np.savez_compressed(path, **out)
    print(f"{name}: rays={n} loss={float(loss):.6e} |gpresc|max={np.abs(gpresc).max():.3e} "
          f"nan_gin={int(np.isnan(gin).sum())} -> {os.path.getsize(path) / 1024:.0f} KiB")


def ref_three_mirror():
    """BASELINE config C3: off-axis three-mirror anastigmat, 4 x 4 polynomial freeform mirrors, rx tilts, dy decenters."""
    from optiland import optic as ref_optic
    from optiland_pr_b200 import samples as my_samples
    to_arr = (lambda a: np.array(a)) if be.get_backend() == "numpy" else (lambda a: [list(r) for r in a])
    return my_samples.build_three_mirror(ref_optic.Optic(name="TMA"), to_array=to_arr)


def ref_freeform_plate(zt):
    """Zernike freeform + odd asphere + tilts about three axes + aperture programs + absorbing glass (the second half
    of BASELINE config C3); the pupil is kept inside the TIR rim so the summed loss is finite."""
    def make():
        from optiland import optic as ref_optic
        import optiland.physical_apertures as ref_ap
        from optiland_pr_b200 import samples as my_samples
        return my_samples.build_freeform_plate(ref_optic.Optic(name="plate"), ref_ap, zt)
    return make


def ref_tilted_singlet():
    """An aspheric singlet whose two surfaces are decentred and tilted about all three axes (refraction through
    tilted frames; the tilt variables of optimization/variable/tilt.py)."""
    lens = AsphericSinglet()
    s1, s2 = lens.surface_group.surfaces[1], lens.surface_group.surfaces[2]
    s1.geometry.cs.rx, s1.geometry.cs.ry, s1.geometry.cs.rz = 0.03, -0.02, 0.4
    s1.geometry.cs.x, s1.geometry.cs.y = 0.15, -0.1
    s2.geometry.cs.rx, s2.geometry.cs.ry = -0.025, 0.015
    s2.geometry.cs.y = 0.2
    return lens


def main():
    only = sys.argv[1:]
    if only:
        cases = {"freeform_plate_fringe": lambda: run_case("freeform_plate_fringe", ref_freeform_plate("fringe"), [0.55], 64, 9, 0.55, False),
                 "freeform_plate_noll": lambda: run_case("freeform_plate_noll", ref_freeform_plate("noll"), [0.55], 64, 10, 0.55, False),
                 "three_mirror": lambda: run_case("three_mirror", ref_three_mirror, [0.587], 48, 7),
                 "tilted_singlet": lambda: run_case("tilted_singlet", ref_tilted_singlet, [0.587], 96, 8)}
        for nm in only:
            cases[nm]()
        return
    run_case("freeform_plate_fringe", ref_freeform_plate("fringe"), [0.55], 64, 9, 0.55, False)
    run_case("freeform_plate_noll", ref_freeform_plate("noll"), [0.55], 64, 10, 0.55, False)
    run_case("three_mirror", ref_three_mirror, [0.587], 48, 7)
    run_case("tilted_singlet", ref_tilted_singlet, [0.587], 96, 8)
    run_case("cooke_triplet", CookeTriplet, [0.48, 0.55, 0.65], 24, 3)
    run_case("double_gauss", DoubleGauss, [0.5876], 64, 4)
    run_case("even_asphere_double_gauss", ref_even_asphere_double_gauss, [0.5876], 64, 5)
    run_case("aspheric_singlet", AsphericSinglet, [0.587], 96, 6)


if __name__ == "__main__":
    main()
