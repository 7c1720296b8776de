"""Spot epilogue: per-(field, wavelength) centroid / RMS / geometric radius from device-side sums.

    RayOperand.rms_spot_size                 optiland/optimization/operand/ray.py:299-340
    SpotDiagram.centroid / rms_spot_radius / geometric_spot_radius
                                             optiland/analysis/spot_diagram.py:317-357

The reductions run in `olrt_spot_sums_*` (csrc/olrt_api.cu: registers -> warp shuffles -> one set
of atomics per CTA); this module only turns the ten sums per group into the scalars the
reference computes with `be.mean` / `be.sqrt` / `be.max` over whole arrays.
"""
from __future__ import annotations

import torch

from . import _lib as L


def spot_sums(x: torch.Tensor, y: torch.Tensor, intensity: torch.Tensor, n_groups: int = 1,
              rays_per_group: int | None = None, shift: torch.Tensor | None = None,
              valid_only: bool = False, out: torch.Tensor | None = None) -> torch.Tensor:
    """[n_groups, 10] float64 device tensor, layout in include/olrt.h (olrt_spot_sums_*).
    Ray r belongs to group r // rays_per_group. Asynchronous on the current stream."""
    lib = L.load()
    if not x.is_cuda:
        raise L.OlrtLibraryError("spot_sums needs CUDA tensors; there is no CPU path")
    n = int(x.numel())
    if rays_per_group is None:
        rays_per_group = (n + n_groups - 1) // n_groups if n else 1
    x, y, intensity = x.contiguous(), y.contiguous(), intensity.contiguous()
    if out is None:
        out = torch.zeros((n_groups, L.SPOT_SUMS), dtype=torch.float64, device=x.device)
    else:
        out.zero_()
    if shift is not None:
        shift = shift.to(device=x.device, dtype=torch.float64).contiguous()
    fn = lib.olrt_spot_sums_f64 if x.dtype == torch.float64 else lib.olrt_spot_sums_f32
    with torch.cuda.device(x.device):
        rc = fn(x.data_ptr(), y.data_ptr(), intensity.data_ptr(), n, int(rays_per_group), int(n_groups),
                0 if shift is None else shift.data_ptr(), L.SPOT_VALID_ONLY if valid_only else 0,
                out.data_ptr(), torch.cuda.current_stream(x.device).cuda_stream)
    L.check(rc, "olrt_spot_sums")
    return out


def spot_stats(sums: torch.Tensor) -> dict:
    """count, centroid and RMS radius about the centroid per group, from one pass of sums taken
    about any shift: mean((x-cx)^2 + (y-cy)^2) = (Sxx + Syy)/n - mx^2 - my^2 with mx = Sx/n."""
    cnt = sums[:, 0]
    mx, my = sums[:, 1] / cnt, sums[:, 2] / cnt
    var = (sums[:, 3] + sums[:, 4]) / cnt - mx * mx - my * my
    return dict(count=cnt, cx=sums[:, 8] + mx, cy=sums[:, 9] + my,
                rms=torch.sqrt(torch.clamp(var, min=0.0)), intensity=sums[:, 6])


def rms_spot_size(x, y, intensity, n_groups=1, rays_per_group=None, valid_only=False):
    """sqrt(mean((x - mean x)^2 + (y - mean y)^2)) per group (ray.py:336-340 when valid_only is
    False: vignetted and NaN rays are included, exactly as the reference's operand does)."""
    return spot_stats(spot_sums(x, y, intensity, n_groups, rays_per_group, None, valid_only))["rms"]


def spot_radii(x, y, intensity, n_groups=1, rays_per_group=None, reference_centroid=None):
    """SpotDiagram-style numbers (valid rays only): centroid, RMS radius and geometric radius about
    `reference_centroid` ([G, 2]; default: each group's own centroid). Two passes."""
    s1 = spot_sums(x, y, intensity, n_groups, rays_per_group, None, True)
    st = spot_stats(s1)
    c = reference_centroid if reference_centroid is not None else torch.stack([st["cx"], st["cy"]], dim=1)
    s2 = spot_sums(x, y, intensity, n_groups, rays_per_group, c, True)
    rms = torch.sqrt((s2[:, 3] + s2[:, 4]) / s2[:, 0])
    geo = torch.sqrt(s2[:, 5])  # stored as the bits of a double (u64-ordered atomicMax)
    return dict(cx=c[:, 0], cy=c[:, 1], rms=rms, geo=geo, count=s2[:, 0])
