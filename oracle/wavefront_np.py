"""TEST INFRASTRUCTURE ONLY -- NumPy restatement of the reference's chief-ray wavefront / OPD epilogue:

    ReferenceStrategy._opd_image_to_xp                 optiland/wavefront/strategy.py:68-116
    ReferenceStrategy._correct_tilt                    optiland/wavefront/strategy.py:118-166
    ChiefRayStrategy.compute_wavefront_data            optiland/wavefront/strategy.py:176-229
    ChiefRayStrategy._calculate_sphere_from_chief_ray  optiland/wavefront/strategy.py:231-240

Inputs are ray states AT THE IMAGE SURFACE (x y z L M N opd intensity), i.e. outputs of the trace; the operation
order follows the reference line by line, so fp64 results are the reference's bits
(tests/test_oracle_golden.py::test_wavefront_oracle_bit_exact).
"""
from __future__ import annotations

import numpy as np


def opd_image_to_xp(x, y, z, L, M, N, xc, yc, zc, R, n_image):
    L, M, N = -L, -M, -N
    a = L**2 + M**2 + N**2
    b = 2 * (L * (x - xc) + M * (y - yc) + N * (z - zc))
    c = x**2 + y**2 + z**2 - 2 * (x * xc + y * yc + z * zc) + xc**2 + yc**2 + zc**2 - R**2
    d = b**2 - 4 * a * c
    d = np.where(d < 0, 0, d)
    t = (-b - np.sqrt(d)) / (2 * a)
    t = np.where(t < 0, (-b + np.sqrt(d)) / (2 * a), t)
    return n_image * t


def tilt_direction(hx, hy, max_field_deg):
    fx_rad, fy_rad = np.deg2rad(hx * max_field_deg), np.deg2rad(hy * max_field_deg)
    tx, ty = np.tan(fx_rad), np.tan(fy_rad)
    uz = 1.0 / np.sqrt(1.0 + tx**2 + ty**2)
    return tx * uz, ty * uz


def correct_tilt(opd, ux, uy, px, py, epd):
    return opd + (ux * (px * epd / 2) + uy * (py * epd / 2))


def chief_ray_wavefront(rays, chief, field, wavelength, px, py, *, epd, max_field, pupil_z, n_image, angle_field=True):
    """rays / chief: dicts of x y z L M N opd (+ i for rays) at the image surface (chief: one ray)."""
    xc, yc, zc = chief["x"], chief["y"], chief["z"]
    with np.errstate(all="ignore"):
        R = float(np.sqrt(xc**2 + yc**2 + (zc - pupil_z) ** 2).item())
        ux, uy = tilt_direction(field[0], field[1], max_field) if angle_field else (0.0, 0.0)
        opd_img_ref = opd_image_to_xp(chief["x"], chief["y"], chief["z"], chief["L"], chief["M"], chief["N"],
                                      xc, yc, zc, R, n_image)
        opd_ref = chief["opd"] - opd_img_ref
        if angle_field:
            opd_ref = correct_tilt(opd_ref, ux, uy, 0, 0, epd)
        opd_img = opd_image_to_xp(rays["x"], rays["y"], rays["z"], rays["L"], rays["M"], rays["N"], xc, yc, zc, R,
                                  n_image)
        opd = rays["opd"] - opd_img
        if angle_field:
            opd = correct_tilt(opd, ux, uy, px, py, epd)
        opd_wv = (opd_ref - opd) / (wavelength * 1e-3)
        t = opd_img / n_image
        return dict(pupil_x=rays["x"] - t * rays["L"], pupil_y=rays["y"] - t * rays["M"],
                    pupil_z=rays["z"] - t * rays["N"], opd=opd_wv, intensity=rays["i"], radius=R)
