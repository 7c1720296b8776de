"""Chief-ray wavefront / OPD analysis on the OPD epilogue kernel.

    Wavefront.get_data                                  optiland/wavefront/wavefront.py:84-100
    ChiefRayStrategy.compute_wavefront_data             optiland/wavefront/strategy.py:176-229
    ReferenceStrategy._opd_image_to_xp / _correct_tilt  optiland/wavefront/strategy.py:68-166
    RayOperand.OPD_difference (uniform weights)         optiland/optimization/operand/ray.py:342-389

Per-ray arithmetic is in `olrt_opd_epilogue_*` (csrc/trace_strict.cu); this module traces the chief ray and the
pupil grid through the fused trace and hands the kernel the per-field scalars (reference sphere, chief-ray OPD, tilt
direction). The paraxial exit-pupil location is an input (`Optic.set_paraxial(EPL, EPD, XPL)`).
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass

import numpy as np
import torch

from . import _lib as L
from . import backend as be
from .optic import pupil_points


@dataclass
class WavefrontData:
    """wavefront/wavefront_data.py:16-34."""
    pupil_x: torch.Tensor
    pupil_y: torch.Tensor
    pupil_z: torch.Tensor
    opd: torch.Tensor          # waves
    intensity: torch.Tensor
    radius: float


def opd_epilogue(rays, Px, Py, sphere: "L.OlrtRefSphere", raw=False, want_pupil=True):
    """Launch `olrt_opd_epilogue_*` on ray states at the image surface. Returns (opd [n], pupil [3, n] or None)."""
    lib = L.load()
    x = rays.x
    if not x.is_cuda:
        raise L.OlrtLibraryError("opd_epilogue needs CUDA tensors; there is no CPU path")
    n, dt, dev = int(x.numel()), x.dtype, x.device
    comps = [rays.x, rays.y, rays.z, rays.L, rays.M, rays.N, rays.i, rays.opd]
    comps = [t.contiguous() for t in comps]
    opd = torch.empty(n, dtype=dt, device=dev)
    pupil = torch.empty((3, n), dtype=dt, device=dev) if want_pupil else None
    if Px is not None:
        Px = Px.to(device=dev, dtype=torch.float64).contiguous()
        Py = Py.to(device=dev, dtype=torch.float64).contiguous()
        if Px.numel() != n:
            raise ValueError("one pupil point per ray is required")
    fn = lib.olrt_opd_epilogue_f64 if dt == torch.float64 else lib.olrt_opd_epilogue_f32
    with torch.cuda.device(dev):
        rc = fn(L.PtrArray8(*[t.data_ptr() for t in comps]), 0 if Px is None else Px.data_ptr(),
                0 if Py is None else Py.data_ptr(), n, C.byref(sphere), opd.data_ptr(),
                0 if pupil is None else pupil.data_ptr(), L.OPD_RAW if raw else 0,
                torch.cuda.current_stream(dev).cuda_stream)
    L.check(rc, "olrt_opd_epilogue")
    return opd, pupil


class ChiefRayStrategy:
    def __init__(self, optic, Px, Py):
        if optic._XPL is None:
            raise ValueError("call set_paraxial(EPL, EPD, XPL) first: the exit-pupil location is an input")
        self.optic = optic
        dev = be.get_device()
        self.Px = torch.as_tensor(np.asarray(Px, dtype=np.float64), device=dev)
        self.Py = torch.as_tensor(np.asarray(Py, dtype=np.float64), device=dev)
        sg = optic.surface_group
        self.pupil_z = optic._XPL + float(sg.surfaces[-1].geometry.cs.z)           # strategy.py:174
        last = sg.surfaces[-2]
        self.n_image = float(last.material_post.n(optic.primary_wavelength))       # optic.n()[-1]

    def compute_wavefront_data(self, field, wavelength) -> WavefrontData:
        optic = self.optic
        hx, hy = float(field[0]), float(field[1])
        keep = optic.surface_group.record
        optic.surface_group.record = "none"
        try:
            chief = optic.trace_generic(hx, hy, 0.0, 0.0, wavelength)
            xc, yc, zc = (float(chief.x[0]), float(chief.y[0]), float(chief.z[0]))
            R = math.sqrt(xc**2 + yc**2 + (zc - self.pupil_z) ** 2)                # strategy.py:231-240
            mf = optic.max_field
            tx, ty = math.tan(math.radians(hx * mf)), math.tan(math.radians(hy * mf))
            uz = 1.0 / math.sqrt(1.0 + tx**2 + ty**2)
            sph = L.OlrtRefSphere(xc, yc, zc, R, self.n_image, 0.0, float(wavelength), tx * uz, ty * uz,
                                  float(optic._EPD))
            opd_ref, _ = opd_epilogue(chief, None, None, sph, raw=True, want_pupil=False)
            sph.opd_ref = float(opd_ref[0])
            rays = optic.trace([hx], [hy], wavelength, distribution=(self.Px, self.Py))
        finally:
            optic.surface_group.record = keep
        opd, pupil = opd_epilogue(rays, self.Px, self.Py, sph)
        return WavefrontData(pupil[0], pupil[1], pupil[2], opd, rays.i, R)


class Wavefront:
    """wavefront/wavefront.py:28-100 for the chief-ray strategy."""

    def __init__(self, optic, fields="all", wavelengths="all", num_rays=12, distribution="hexapolar", seed=None):
        self.optic = optic
        self.fields = optic.get_field_coords() if isinstance(fields, str) else list(fields)
        self.wavelengths = list(optic.wavelengths) if isinstance(wavelengths, str) else list(wavelengths)
        Px, Py = pupil_points(distribution, num_rays, seed) if isinstance(distribution, str) else distribution
        self.strategy = ChiefRayStrategy(optic, Px, Py)
        self.data = {}

    def get_data(self, field, wl) -> WavefrontData:
        key = (tuple(field), float(wl))
        if key not in self.data:
            self.data[key] = self.strategy.compute_wavefront_data(field, wl)
        return self.data[key]


def opd_difference(optic, Hx, Hy, num_rays, wavelength, distribution="hexapolar"):
    """mean |opd - mean(opd)| over the pupil grid (operand/ray.py:342-389 with unit weights)."""
    d = Wavefront(optic, [(Hx, Hy)], [wavelength], num_rays, distribution).get_data((Hx, Hy), wavelength)
    return torch.mean(torch.abs(d.opd - torch.mean(d.opd)))
