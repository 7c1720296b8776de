"""Shared helpers for the parity tests."""
import os

import numpy as np

from optiland_pr_b200 import samples
from optiland_pr_b200.prescription import pack_surface_group

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
FIELDS8 = ("x", "y", "z", "L", "M", "N", "intensity", "opd")

MIRRORS = {
    "cooke_triplet": samples.CookeTriplet,
    "double_gauss": samples.DoubleGauss,
    "even_asphere_double_gauss": samples.EvenAsphereDoubleGauss,
    "aspheric_singlet": samples.AsphericSinglet,
    "three_mirror": samples.ThreeMirrorAnastigmat,
    "freeform_plate_fringe": lambda: samples.FreeformPlate("fringe"),
    "freeform_plate_noll": lambda: samples.FreeformPlate("noll"),
    "freeform_plate_standard": lambda: samples.FreeformPlate("standard"),
    "anamorphic": samples.Anamorphic,
    "chebyshev_plate": samples.ChebyshevPlate,
    "coated_relay": samples.CoatedRelay,
}
# lenses with a gradient fixture only (tests/golden/<name>_grads.npz)
GRAD_ONLY_MIRRORS = {"tilted_singlet": samples.TiltedSinglet}
# lenses inside the backward kernel's differentiable set (gradient fixtures exist for these)
GRAD_MIRRORS = ("cooke_triplet", "double_gauss", "even_asphere_double_gauss", "aspheric_singlet")
# lenses of the general adjoint kernel only: tilted frames, x^i y^j polynomial freeforms (tilt and polynomial-coefficient
# variables of the reference: optimization/variable/tilt.py:59, polynomial_coeff.py:72)
GRAD_MIRRORS_GENERAL = ("three_mirror", "tilted_singlet", "freeform_plate_fringe", "freeform_plate_noll")
# fixtures whose surfaces are all closed-form (+,-,*,/,sqrt only): oracle must match the reference bit for bit
CLOSED_FORM = ("cooke_triplet", "double_gauss", "coated_relay")


def load_golden(name):
    with np.load(os.path.join(GOLDEN, name + ".npz")) as z:
        return {k: z[k] for k in z.files}


def mirror_packed(name, wavelengths, record="all"):
    lens = (MIRRORS.get(name) or GRAD_ONLY_MIRRORS[name])()
    return lens, pack_surface_group(lens.surface_group.surfaces, list(wavelengths), record=record)


def rel_err(a, b):
    """max |a-b| / max(|b|, 1) over finite entries; NaN patterns must agree."""
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    assert np.array_equal(np.isnan(a), np.isnan(b)), "NaN pattern differs"
    m = ~np.isnan(b)
    if not m.any():
        return 0.0
    return float(np.max(np.abs(a[m] - b[m]) / np.maximum(np.abs(b[m]), 1.0)))


def fp32_errors(rec, ref):
    """fp32 parity numbers (DESIGN.md "fp32 parity"); rec, ref: [S, 8, N].
    positions : max |dp| over the rays of a surface / the largest |coordinate| of the bundle there
                (relative error of the position vectors), worst surface;
    directions: max |dL|, |dM|, |dN| (unit vectors: absolute = relative);
    opd       : relative to the OPD itself."""
    rec, ref = np.asarray(rec, dtype=np.float64), np.asarray(ref, dtype=np.float64)
    assert np.array_equal(np.isnan(rec), np.isnan(ref)), "NaN pattern differs"
    pos = dirs = opd = 0.0
    for s in range(ref.shape[0]):
        m = ~np.isnan(ref[s]).any(axis=0)
        if not m.any():
            continue
        scale = max(float(np.max(np.abs(ref[s, 0:3][:, m]))), 1.0)
        pos = max(pos, float(np.max(np.abs(rec[s, 0:3][:, m] - ref[s, 0:3][:, m]))) / scale)
        dirs = max(dirs, float(np.max(np.abs(rec[s, 3:6][:, m] - ref[s, 3:6][:, m]))))
        opd = max(opd, float(np.max(np.abs(rec[s, 7][m] - ref[s, 7][m]) / np.maximum(np.abs(ref[s, 7][m]), 1.0))))
    return pos, dirs, opd
