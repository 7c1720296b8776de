"""TEST INFRASTRUCTURE ONLY -- tests/golden/wavefront_cooke.npz from the UNMODIFIED reference:
Wavefront(optic, ..., strategy="chief_ray").get_data(field, wl) for the Cooke triplet (wavefront/wavefront.py,
wavefront/strategy.py:176-229), together with everything needed to recompute it without the reference: the traced
ray states at the image, the chief ray, the pupil grid and the paraxial numbers.   python oracle/make_golden_wavefront.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_shim  # noqa: E402

ref_shim.install()
import optiland.backend as be  # noqa: E402
from optiland.samples.objectives import CookeTriplet  # noqa: E402
from optiland.wavefront import Wavefront  # noqa: E402


def main():
    be.set_backend("numpy")
    lens = CookeTriplet()
    fields = [(0.0, 0.0), (0.0, 0.7), (0.0, 1.0)]
    wl = 0.55
    wf = Wavefront(lens, fields, [wl], num_rays=8, distribution="hexapolar", strategy="chief_ray")
    strat = wf.strategy
    out = dict(fields=np.array(fields), wavelength=wl, px=np.asarray(strat.distribution.x), py=np.asarray(strat.distribution.y),
               EPD=float(lens.paraxial.EPD()), EPL=float(lens.paraxial.EPL()), XPL=float(lens.paraxial.XPL()),
               pupil_z=float(strat.pupil_z), n_image=float(strat.n_image), max_field=float(lens.fields.max_field))
    names = ("x", "y", "z", "L", "M", "N", "opd")
    for k, f in enumerate(fields):
        d = wf.get_data(f, wl)
        out[f"f{k}_pupil"] = np.stack([d.pupil_x, d.pupil_y, d.pupil_z])
        out[f"f{k}_opd"] = np.asarray(d.opd)
        out[f"f{k}_intensity"] = np.asarray(d.intensity)
        out[f"f{k}_radius"] = float(d.radius)
        chief = lens.trace_generic(*f, Px=0.0, Py=0.0, wavelength=wl)
        out[f"f{k}_chief"] = np.stack([np.asarray(getattr(chief, n_), dtype=np.float64).reshape(-1) for n_ in names])
        rays = lens.trace(*f, wl, None, strat.distribution)
        out[f"f{k}_rays"] = np.stack([np.asarray(getattr(rays, n_), dtype=np.float64) for n_ in names] +
                                     [np.asarray(lens.surface_group.intensity[-1, :], dtype=np.float64)])
    path = os.path.join(ROOT, "tests", "golden", "wavefront_cooke.npz")
    np.savez_compressed(path, **out)
    print(path, {k: getattr(v, "shape", v) for k, v in out.items()})


if __name__ == "__main__":
    main()
