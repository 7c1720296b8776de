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
}
# fixtures whose surfaces are all closed-form (+,-,*,/,sqrt only): oracle must match the reference bit for bit
CLOSED_FORM = ("cooke_triplet", "double_gauss")


def load_golden(name):
    with np.load(os.path.join(GOLDEN, name + ".npz")) as z:
        return {k: z[k] for k in z.files}


def mirror_packed(name, wavelengths, record="all"):
    lens = MIRRORS[name]()
    return lens, pack_surface_group(lens.surface_group.surfaces, list(wavelengths), record=record)


def rel_err(a, b):
    """max |a-b| / max(|b|, 1) over finite entries; NaN patterns must agree."""
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    assert np.array_equal(np.isnan(a), np.isnan(b)), "NaN pattern differs"
    m = ~np.isnan(b)
    if not m.any():
        return 0.0
    return float(np.max(np.abs(a[m] - b[m]) / np.maximum(np.abs(b[m]), 1.0)))


def scale_rel_err(a, b):
    """fp32 tolerance (DESIGN.md "fp32 parity"): per surface and component, max |a-b| over the rays
    divided by the scale of that component on that surface, max(max |b|, 1). a, b: [S, N]."""
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    assert np.array_equal(np.isnan(a), np.isnan(b)), "NaN pattern differs"
    worst = 0.0
    for s in range(b.shape[0]):
        m = ~np.isnan(b[s])
        if m.any():
            worst = max(worst, float(np.max(np.abs(a[s][m] - b[s][m])) / max(float(np.max(np.abs(b[s][m]))), 1.0)))
    return worst
