"""Workloads run through the UNMODIFIED reference's own classes (imported via oracle/ref_shim.py), shared by the
drop-in tests: the same function is called once on the reference's stock path (NumPy backend, or torch-CPU fp64 for
gradients = the oracle of record) and once on torch + CUDA with `optiland_pr_b200.dropin.install()`.
Every function returns plain NumPy / float data."""
import numpy as np

from oracle import ref_shim

ref_shim.install()

import optiland.backend as be  # noqa: E402


def to_np(v):
    if hasattr(v, "detach"):
        return v.detach().cpu().numpy().astype(np.float64)
    return np.asarray(v, dtype=np.float64)


def use_numpy():
    be.set_backend("numpy")


def use_torch(device, precision="float64", grad=False):
    """`torch.set_default_device` goes with `be.set_device`: the unmodified reference creates some tensors without
    its configured device (`be.atleast_1d`, torch_backend.py:754-756, on the Hx / Hy of every Optic.trace), so its
    own CUDA path only runs when torch's default device is the GPU as well. Configuration, not a patch."""
    import torch
    be.set_backend("torch")
    be.set_device(device)
    torch.set_default_device(device)
    be.set_precision(precision)
    if grad:
        be.grad_mode.enable()
    else:
        be.grad_mode.disable()


def restore():
    """Leave the reference's globals as a fresh import has them."""
    try:
        import torch
        torch.set_default_device("cpu")
        be.set_backend("torch")
        be.set_device("cpu")
        be.set_precision("float32")
        be.grad_mode.disable()
    finally:
        be.set_backend("numpy")


def make_lens(name):
    from optiland.samples import objectives, simple, telescopes
    from optiland.physical_apertures import RadialAperture
    if name == "cooke":
        return objectives.CookeTriplet()
    if name == "double_gauss":
        return objectives.DoubleGauss()
    if name == "hubble":
        return telescopes.HubbleTelescope()
    if name == "aspheric_singlet":
        return simple.AsphericSinglet()
    if name == "reverse_telephoto":
        return objectives.ReverseTelephoto()
    if name == "cooke_stopped":   # physical apertures that vignette part of the off-axis bundle
        lens = objectives.CookeTriplet()
        lens.surface_group.surfaces[1].aperture = RadialAperture(r_max=4.6)
        lens.surface_group.surfaces[6].aperture = RadialAperture(r_max=4.0, r_min=0.3)
        return lens
    raise KeyError(name)


def spot_diagram(name):
    """SpotDiagram (analysis/spot_diagram.py:397-438): rms / geometric radii [field][wavelength] + centroids."""
    from optiland.analysis import SpotDiagram
    spot = SpotDiagram(make_lens(name))
    return {"rms": np.array([[float(to_np(v)) for v in row] for row in spot.rms_spot_radius()]),
            "geo": np.array([[float(to_np(v)) for v in row] for row in spot.geometric_spot_radius()]),
            "centroid": np.array([[float(to_np(c)) for c in row] for row in spot.centroid()])}


def rms_spot_operand(name, Hx, Hy, wavelength, num_rays, distribution="hexapolar"):
    """RayOperand.rms_spot_size (optimization/operand/ray.py:299-340)."""
    from optiland.optimization.operand.ray import RayOperand
    v = RayOperand.rms_spot_size(optic=make_lens(name), surface_number=-1, Hx=Hx, Hy=Hy, wavelength=wavelength,
                                 num_rays=num_rays, distribution=distribution)
    return float(to_np(v))


def ray_operands(name):
    """Single-ray operands reading positive / negative surface indices (operand/ray.py:57, :262-264)."""
    from optiland.optimization.operand.ray import RayOperand
    lens = make_lens(name)
    kw = dict(optic=lens, Hx=0.0, Hy=1.0, Px=0.3, Py=-0.6, wavelength=0.55)
    out = {}
    for s in (1, 3, -1):
        for nm in ("x_intercept", "y_intercept", "z_intercept", "L", "M", "N"):
            out[f"{nm}[{s}]"] = float(to_np(getattr(RayOperand, nm)(surface_number=s, **kw)))
    return out


def opd_difference(name, Hx, Hy, num_rays, distribution="gaussian_quad"):
    """RayOperand.OPD_difference (operand/ray.py:342-389)."""
    from optiland.optimization.operand.ray import RayOperand
    return float(to_np(RayOperand.OPD_difference(optic=make_lens(name), Hx=Hx, Hy=Hy, wavelength=0.55,
                                                 num_rays=num_rays, distribution=distribution)))


def wavefront(name, field=(0.0, 1.0), wl=0.55, num_rays=12):
    """Wavefront(...).get_data (wavefront/strategy.py:176-229): OPD in waves, pupil points, intensity, radius."""
    from optiland.wavefront import Wavefront
    wf = Wavefront(make_lens(name), fields=[field], wavelengths=[wl], num_rays=num_rays, distribution="hexapolar")
    d = wf.get_data(field, wl)
    return {"opd": to_np(d.opd), "pupil_x": to_np(d.pupil_x), "pupil_y": to_np(d.pupil_y), "pupil_z": to_np(d.pupil_z),
            "intensity": to_np(d.intensity), "radius": float(to_np(d.radius))}


def trace_records(name, Hx, Hy, wl, num_rays, distribution):
    """Optic.trace, then every per-surface array of surface_group (surface_group.py:95-140) + the returned rays."""
    lens = make_lens(name)
    rays = lens.trace(Hx, Hy, wl, num_rays, distribution)
    sg = lens.surface_group
    out = {k: to_np(getattr(sg, k)) for k in ("x", "y", "z", "L", "M", "N", "intensity", "opd")}
    for k in ("x", "y", "z", "L", "M", "N", "i", "opd", "L0", "M0", "N0"):
        out["rays." + k] = to_np(getattr(rays, k))
    return out


def trace_generic(name, Hx, Hy, Px, Py, wl):
    """Optic.trace_generic (raytrace/real_ray_tracer.py:84-123), scalar or array arguments."""
    lens = make_lens(name)
    if not np.isscalar(Hx):
        Hx, Hy, Px, Py = (be.array(np.asarray(v, dtype=np.float64)) for v in (Hx, Hy, Px, Py))
    rays = lens.trace_generic(Hx, Hy, Px, Py, wl)
    sg = lens.surface_group
    out = {k: to_np(getattr(sg, k)) for k in ("x", "y", "z", "L", "M", "N", "intensity", "opd")}
    out["rays.y"] = to_np(rays.y)
    return out


def adam_run(name, n_steps=4, lr=None, num_rays=6):
    """TorchAdamOptimizer (optimization/optimizer/torch/base.py:95-154) on RMS-spot operands (3 fields) with radius,
    thickness (and conic / asphere-coefficient) variables. Returns per-step losses, per-step parameter gradients and
    the final parameter values.
    Two facts about the unmodified reference shape this harness (both arms get the same treatment):
      * its material objects memoise n(w) / k(w) tensors WITH their autograd graph (materials/base.py:73-119), so a
        second optimiser step fails with "backward through the graph a second time": the per-step callback clears
        those caches;
      * `set_thickness` rebuilds every cs.z from detached positions (optic_updater.py:67-85), so of several thickness
        variables only the last keeps a gradient: one thickness variable per lens is used."""
    from optiland.optimization import OptimizationProblem
    from optiland.optimization.optimizer.torch import TorchAdamOptimizer
    lens = make_lens(name)
    problem = OptimizationProblem()
    for Hy in (0.0, 0.7, 1.0):
        problem.add_operand(operand_type="rms_spot_size", target=0.0, weight=1.0,
                            input_data={"optic": lens, "surface_number": -1, "Hx": 0.0, "Hy": Hy, "num_rays": num_rays,
                                        "wavelength": 0.55, "distribution": "hexapolar"})
    if name == "cooke":
        for s in (1, 2, 3, 4, 5, 6):
            problem.add_variable(lens, "radius", surface_number=s)
        problem.add_variable(lens, "thickness", surface_number=5)
        lr = 1e-3 if lr is None else lr
    elif name == "aspheric_singlet":
        problem.add_variable(lens, "radius", surface_number=1)
        problem.add_variable(lens, "conic", surface_number=1)
        for k in range(3):
            problem.add_variable(lens, "asphere_coeff", surface_number=1, coeff_number=k)
        problem.add_variable(lens, "thickness", surface_number=2)
        lr = 1e-8 if lr is None else lr      # Adam moves every (scaled) parameter by ~lr per step
    else:
        raise KeyError(name)
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        opt = TorchAdamOptimizer(problem)
        losses = []
        grads = []

        def cb(i, loss):
            losses.append(float(loss))
            grads.append([float("nan") if p.grad is None else float(p.grad.detach().cpu()) for p in opt.params])
            for s in lens.surface_group.surfaces:
                for m in (s.material_pre, s.material_post):
                    if m is not None:
                        m._n_cache.clear()
                        m._k_cache.clear()

        res = opt.optimize(n_steps=n_steps, lr=lr, gamma=0.99, disp=False, callback=cb)
    return {"losses": np.array(losses), "grads": np.array(grads), "x": np.array([float(v) for v in res.x]),
            "final": float(res.fun)}
