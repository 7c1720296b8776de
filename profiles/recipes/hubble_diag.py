"""Diagnostic: per-surface deviation of the drop-in (tolerance / strict) from the NumPy reference on the Hubble."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import ref_cases as rc
from optiland_pr_b200 import dropin

name = sys.argv[1] if len(sys.argv) > 1 else "hubble"
args = (0.0, 1.0, 0.55, 24, "hexapolar")
rc.use_numpy()
want = rc.trace_records(name, *args)
res = {}
for mode, fast in (("tolerance", True), ("strict", False)):
    rc.use_torch("cuda", "float64")
    dropin.install(fast_math=fast)
    res[mode] = rc.trace_records(name, *args)
    dropin.uninstall()
rc.use_torch("cuda", "float64")
res["eager"] = rc.trace_records(name, *args)
np.set_printoptions(linewidth=200, precision=2)
for k in ("x", "y", "z", "L", "M", "N", "opd"):
    for mode in res:
        d = np.nanmax(np.abs(res[mode][k] - want[k]), axis=1)
        print(f"{k:4s} {mode:10s} max|d| per surface {d}   scale {np.nanmax(np.abs(want[k]), axis=1)}")
