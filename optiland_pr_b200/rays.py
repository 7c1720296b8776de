"""RealRays container mirroring optiland/rays/real_rays.py:21-88.

Only the container: rotate / refract / reflect / clip (real_rays.py:90-181) are executed inside
the fused CUDA kernel (csrc/trace_common.cuh), not as tensor passes here.
"""
from __future__ import annotations

import torch

from . import backend as be


class BaseRays:
    """optiland/rays/base.py:18-42"""


class RealRays(BaseRays):
    """x, y, z, L, M, N, i (intensity), w (wavelength, um), opd as 1-D CUDA tensors."""

    def __init__(self, x, y, z, L, M, N, intensity, wavelength):
        self.x = be.as_array_1d(x)
        self.y = be.as_array_1d(y)
        self.z = be.as_array_1d(z)
        self.L = be.as_array_1d(L)
        self.M = be.as_array_1d(M)
        self.N = be.as_array_1d(N)
        self.i = be.as_array_1d(intensity)
        self.w = be.as_array_1d(wavelength)
        n = max(t.numel() for t in (self.x, self.y, self.z, self.L, self.M, self.N, self.i, self.w))
        for name in ("x", "y", "z", "L", "M", "N", "i", "w"):
            t = getattr(self, name)
            if t.numel() == 1 and n > 1:
                setattr(self, name, t.expand(n).contiguous())
        self.opd = torch.zeros_like(self.x)
        # pre-surface direction cosines (local frame of the last surface), real_rays.py:82-86
        self.L0 = None
        self.M0 = None
        self.N0 = None
        self.is_normalized = True

    @classmethod
    def from_traced(cls, out, dir0, wavelength):
        """Rays whose launch state never existed in memory (the kernel launched them, olrt_trace_pupil_*):
        rows of `out` [8, N] are the state after the last surface, `dir0` [3, N] the pre-interaction direction
        there; the wavelength is a stride-0 view."""
        r = cls.__new__(cls)
        r.x, r.y, r.z, r.L, r.M, r.N, r.i, r.opd = (out[q] for q in range(8))
        r.w = torch.as_tensor(float(wavelength), dtype=out.dtype, device=out.device).expand(out.shape[1])
        r.L0, r.M0, r.N0 = (dir0[0], dir0[1], dir0[2]) if dir0 is not None else (None, None, None)
        r.is_normalized = True
        return r

    def __len__(self):
        return int(self.x.numel())

    def components(self):
        return [self.x, self.y, self.z, self.L, self.M, self.N, self.i, self.w]
