"""TEST INFRASTRUCTURE ONLY -- tests/golden/spot_diagram.npz from the UNMODIFIED reference: SpotDiagram(lens).centroid()
/ rms_spot_radius() / geometric_spot_radius() (analysis/spot_diagram.py:317-357) for all fields x wavelengths of the
Cooke triplet and of the even-asphere double Gauss (hexapolar, 6 rings).     python oracle/make_golden_spot.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_shim  # noqa: E402

ref_shim.install()
import optiland.backend as be  # noqa: E402
from optiland.analysis import SpotDiagram  # noqa: E402
from optiland.samples.objectives import CookeTriplet  # noqa: E402

from oracle.make_golden import ref_even_asphere_double_gauss  # noqa: E402


def main():
    be.set_backend("numpy")
    out = {}
    for name, lens in (("cooke", CookeTriplet()), ("eadg", ref_even_asphere_double_gauss())):
        sd = SpotDiagram(lens, num_rings=6, distribution="hexapolar")
        out[f"{name}_centroid"] = np.array([[float(c[0]), float(c[1])] for c in sd.centroid()])
        out[f"{name}_rms"] = np.array([[float(v) for v in row] for row in sd.rms_spot_radius()])
        out[f"{name}_geo"] = np.array([[float(v) for v in row] for row in sd.geometric_spot_radius()])
        out[f"{name}_wavelengths"] = np.array([float(w.value) for w in lens.wavelengths.wavelengths])
        out[f"{name}_paraxial"] = np.array([float(lens.paraxial.EPL()), float(lens.paraxial.EPD())])
    path = os.path.join(ROOT, "tests", "golden", "spot_diagram.npz")
    np.savez_compressed(path, **out)
    print(path, {k: v for k, v in out.items()})


if __name__ == "__main__":
    main()
