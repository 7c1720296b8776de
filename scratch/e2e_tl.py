"""e2e step timeline (debug_timeline) + raw pinned H2D bandwidth + chunk sweep."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import optiland_pr_b200 as ol
from optiland_pr_b200 import samples, spot
ol.backend.set_device("cuda:0"); ol.backend.set_precision("float64")
npf = 1 << 24
lens = samples.EvenAsphereDoubleGauss(); lens.surface_group.record = "none"
rng = np.random.default_rng(1)
r_ = rng.uniform(size=npf); th = rng.uniform(0, 2*np.pi, size=npf)
Px_h = torch.from_numpy(np.sqrt(r_)*np.cos(th)).pin_memory(); Py_h = torch.from_numpy(np.sqrt(r_)*np.sin(th)).pin_memory()
Hx = [0.0, 0.0, 0.0]; Hy = [0.0, 10/14, 1.0]
d = torch.empty(npf, dtype=torch.float64, device="cuda")
for _ in range(3): d.copy_(Px_h, non_blocking=True)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10): d.copy_(Px_h, non_blocking=True)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
print(f"H2D pinned 134 MB: {ms:.3f} ms = {npf*8/ms/1e6:.1f} GB/s")
rms_host = torch.empty(3, dtype=torch.float64).pin_memory()
def step():
    rr = lens.trace(Hx, Hy, 0.5876, distribution=(Px_h, Py_h), spot=True)
    rms_host.copy_(spot.spot_stats(rr.spot_sums_host)["rms"])
def timed(k=6):
    for _ in range(3): step()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(k): step()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / k * 1e3
ref = None
for st_, c in ((False, 8), (True, 8), (True, 12), (True, 16), (True, 24)):
    lens.pipeline_streamed = st_; lens.pipeline_chunks = c
    print("streamed", st_, "chunks", c, "e2e ms/step %.3f" % timed(), rms_host.tolist())
lens.pipeline_chunks = 8
lens.debug_timeline = True
for _ in range(3):
    t0 = time.perf_counter(); step(); t1 = time.perf_counter()
    print("host ms %.2f" % ((t1 - t0) * 1e3), " ".join(f"{lb}:{t:.2f}" for lb, t in lens.last_timeline))
