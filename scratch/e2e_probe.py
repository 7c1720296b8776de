"""Where does the e2e step go? host-side cProfile of Optic.trace (pipelined) + chunk-count sweep."""
import os, sys, time, cProfile, pstats
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import optiland_pr_b200 as ol
from optiland_pr_b200 import samples, spot, _lib
ol.backend.set_device("cuda:0"); ol.backend.set_precision("float64")
npf = 1 << 24
lens = samples.EvenAsphereDoubleGauss(); lens.surface_group.record = "none"
rng = np.random.default_rng(1)
r_ = rng.uniform(size=npf); th = rng.uniform(0, 2*np.pi, size=npf)
Px_h = torch.from_numpy(np.sqrt(r_)*np.cos(th)).pin_memory(); Py_h = torch.from_numpy(np.sqrt(r_)*np.sin(th)).pin_memory()
Hx = [0.0, 0.0, 0.0]; Hy = [0.0, 10/14, 1.0]
sums = torch.zeros((3, _lib.SPOT_SUMS), dtype=torch.float64, device="cuda")
rms_host = torch.empty(3, dtype=torch.float64).pin_memory()
def step():
    rr = lens.trace(Hx, Hy, 0.5876, distribution=(Px_h, Py_h))
    s = spot.spot_sums(rr.x, rr.y, rr.i, 3, npf, out=sums)
    rms_host.copy_(spot.spot_stats(s)["rms"])
def timed(k=5):
    for _ in range(3): step()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(k): step()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / k * 1e3
for c in (1, 2, 4, 8):
    lens.pipeline_chunks = c
    print("chunks", c, "e2e ms/step %.3f" % timed())
lens.pipeline_chunks = 4
torch.cuda.synchronize()
for _ in range(3):
    t0 = time.perf_counter(); rr = lens.trace(Hx, Hy, 0.5876, distribution=(Px_h, Py_h)); t1 = time.perf_counter()
    ea, eb, ec = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    ea.record()
    s_ = spot.spot_sums(rr.x, rr.y, rr.i, 3, npf, out=sums); t2 = time.perf_counter(); eb.record()
    st_ = spot.spot_stats(s_)["rms"]; t3 = time.perf_counter(); ec.record()
    torch.cuda.current_stream().synchronize(); t3b = time.perf_counter()
    torch.cuda.synchronize(); t3c = time.perf_counter()
    rms_host.copy_(st_); t4 = time.perf_counter()
    print("host ms: stream sync %.2f device sync %.2f" % ((t3b-t3)*1e3, (t3c-t3b)*1e3))
    print("gpu ms: spot_sums %.3f stats %.3f" % (ea.elapsed_time(eb), eb.elapsed_time(ec)))
    print("host ms: trace %.2f spot_sums %.2f spot_stats %.2f copy %.2f" % ((t1-t0)*1e3, (t2-t1)*1e3, (t3-t2)*1e3, (t4-t3)*1e3))
from optiland_pr_b200.trace import fused_trace
packed = lens.surface_group.pack([0.5876])
for nn in (1 << 22, 1 << 24):
    ins = [rr.x[:nn].clone() for _ in range(7)]
    g = torch.Generator(device="cuda"); 
    rays = lens.launch_rays(np.array([0.0]), np.array([0.7]), Px_h[:nn].cuda(), Py_h[:nn].cuda(), 0.5876)
    ins = rays.components(); ins[7] = None
    o = torch.empty((8, nn), dtype=torch.float64, device="cuda")
    for _ in range(3): fused_trace(packed, ins, out=o, fast=True, verify_nr=False, check_status=False)
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(10): fused_trace(packed, ins, out=o, fast=True, verify_nr=False, check_status=False)
    e1.record(); torch.cuda.synchronize()
    print("fx kernel n=%d: %.3f ms/launch" % (nn, e0.elapsed_time(e1) / 10))
lens.debug_timeline = True
step(); step()
print(" ".join("%s=%.2f" % t for t in lens.last_timeline))
lens.debug_timeline = False
pr = cProfile.Profile(); pr.enable()
for _ in range(5): step()
torch.cuda.synchronize(); pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(8)
