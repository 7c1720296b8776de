"""TEST INFRASTRUCTURE ONLY -- NumPy float64 restatement of the reference's ray launch (non-telecentric object
space): angle fields with the object at infinity or at a finite distance, object-height fields, vignetting factors:

    RayGenerator.generate_rays          optiland/rays/ray_generator.py:28-106
    AngleField.get_ray_origins          optiland/fields/field_types.py:139-169
    ObjectHeightField.get_ray_origins   optiland/fields/field_types.py:245-272
    FieldGroup.get_vig_factor           optiland/fields/field_group.py:80-109
    StandardGeometry.sag                optiland/geometries/standard.py:73-87
    AngleField._get_starting_z_offset   optiland/fields/field_types.py:223-239
    RandomDistribution.generate_points  optiland/distribution.py:132-158

EPL / EPD (paraxial, optiland/paraxial.py:207-280) are inputs. Used by tests and by the CPU
baseline leg of bench.py; never by the product path.
"""
from __future__ import annotations

import numpy as np


def random_pupil(num_points: int, seed=None):
    rng = np.random.default_rng(seed)
    r = rng.uniform(size=num_points)
    theta = rng.uniform(0, 2 * np.pi, size=num_points)
    return np.sqrt(r) * np.cos(theta), np.sqrt(r) * np.sin(theta)


def hexapolar_pupil(num_rings: int):
    x, y = np.zeros(1), np.zeros(1)
    for i in range(num_rings):
        num_theta = 6 * (i + 1)
        theta = np.linspace(0, 2 * np.pi, num_theta + 1)[:-1]
        r = (i + 1) / num_rings
        x = np.concatenate([x, r * np.cos(theta)])
        y = np.concatenate([y, r * np.sin(theta)])
    return x, y


def vig_factor(fields_xy, vig_xy, max_field, Hx, Hy):
    """field_group.py:80-109: vignetting factors of the nearest defined field, in normalised coordinates."""
    pts = np.asarray(fields_xy, dtype=np.float64).reshape(-1, 2)
    if max_field != 0:
        pts = pts / max_field
    d2 = (Hx - pts[:, 0]) ** 2 + (Hy - pts[:, 1]) ** 2
    k = int(np.argmin(d2))
    return float(vig_xy[k][0]), float(vig_xy[k][1])


def standard_sag(x, y, radius, conic):
    """standard.py:84-87."""
    r2 = x ** 2 + y ** 2
    return r2 / (radius * (1 + np.sqrt(1 - (1 + conic) * r2 / radius ** 2)))


def generate_rays(Hx, Hy, Px, Py, wavelength, *, EPL, EPD, max_field, surface_z, vx=1.0, vy=1.0,
                  object_z=None, field_type="angle", object_radius=np.inf, object_conic=0.0):
    """surface_z: z position of every surface except the object (positions[1:]); the image is last.
    vx, vy: 1 - vignetting factor of the field. object_z: None for an object at infinity, else positions[0].
    Returns x0 y0 z0 L M N intensity wavelength."""
    Px, Py = np.asarray(Px, dtype=np.float64), np.asarray(Py, dtype=np.float64)
    field_x = max_field * Hx
    field_y = max_field * Hy
    if field_type == "object_height":
        if object_z is None:
            raise ValueError('Field type cannot be "object_height" for an object at infinity.')
        x0, y0 = np.float64(field_x), np.float64(field_y)
        sag = 0.0 if np.isinf(object_radius) else standard_sag(x0, y0, object_radius, object_conic)
        z0 = sag + object_z
        x0, y0, z0 = np.full_like(Px, x0), np.full_like(Px, y0), np.full_like(Px, z0)
    elif object_z is None:
        z_inner = np.asarray(surface_z[:-1], dtype=np.float64)
        offset = EPD - np.min(z_inner)
        x = -np.tan(np.radians(field_x)) * (offset + EPL)
        y = -np.tan(np.radians(field_y)) * (offset + EPL)
        z = surface_z[0] - offset
        x0 = Px * EPD / 2 * vx + x
        y0 = Py * EPD / 2 * vy + y
        z0 = np.full_like(Px, z)
    else:
        x0 = np.full_like(Px, -np.tan(np.radians(field_x)) * (EPL - object_z))
        y0 = np.full_like(Px, -np.tan(np.radians(field_y)) * (EPL - object_z))
        z0 = np.full_like(Px, object_z)
    x1 = Px * EPD * vx / 2
    y1 = Py * EPD * vy / 2
    z1 = np.full_like(Px, EPL)
    mag = np.sqrt((x1 - x0) ** 2 + (y1 - y0) ** 2 + (z1 - z0) ** 2)
    is_zero = mag < 1e-9
    mag = np.where(is_zero, 1.0, mag)
    L = np.where(is_zero, 0.0, (x1 - x0) / mag)
    M = np.where(is_zero, 0.0, (y1 - y0) / mag)
    N = np.where(is_zero, 1.0, (z1 - z0) / mag)
    return x0, y0, z0, L, M, N, np.ones_like(Px), np.ones_like(x1) * wavelength
