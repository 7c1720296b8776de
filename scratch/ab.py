"""A/B timing of the forward kernels on the bench lens. usage: python scratch/ab.py [nrays]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import optiland_pr_b200 as ol
from optiland_pr_b200 import samples
from optiland_pr_b200.trace import fused_trace
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 24
dev = torch.device("cuda")
lens = samples.EvenAsphereDoubleGauss(); lens.surface_group.record = "none"
rng = np.random.default_rng(0)
r = torch.from_numpy(rng.uniform(size=n)).to(dev); th = torch.from_numpy(rng.uniform(0, 2*np.pi, n)).to(dev)
rays = lens.generate_rays(0.0, 0.7, torch.sqrt(r)*torch.cos(th), torch.sqrt(r)*torch.sin(th), 0.5876)
ins = rays.components(); ins[7] = None
packed = lens.surface_group.pack([0.5876])
out = torch.empty((8, n), dtype=torch.float64, device=dev)
fused_trace(packed, ins, out=out)
def t(fn, k=5):
    for _ in range(2): fn()
    torch.cuda.synchronize(); e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(k): fn()
    e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1)/k
ms = t(lambda: fused_trace(packed, ins, out=out, verify_nr=False, check_status=False))
from optiland_pr_b200 import _lib as L
res = torch.zeros((packed.n_surfaces, L.RESID_STRIDE), dtype=torch.int64, device=dev)
msr = t(lambda: fused_trace(packed, ins, out=out, resid_out=res, check_status=False))
msx = t(lambda: fused_trace(packed, ins, out=out, verify_nr=False, fast=True, replay=True, check_status=False))
msf = t(lambda: fused_trace(packed, ins, out=out, verify_nr=False, fast=True, check_status=False))
ins32 = [x.float() if x is not None else None for x in ins]; o32 = torch.empty((8, n), dtype=torch.float32, device=dev)
ms32 = t(lambda: fused_trace(packed, ins32, out=o32, verify_nr=False, check_status=False))
ms32r = t(lambda: fused_trace(packed, ins32, out=o32, verify_nr=False, replay=True, check_status=False))
a = fused_trace(packed, ins).out.clone(); rb = fused_trace(packed, ins, fast=True); b = rb.out
den = torch.clamp(a.abs(), min=1.0)
err = ((a - b).abs() / den).nan_to_num(0.0).amax(dim=1).cpu().numpy()
same_nan = bool((a.isnan() == b.isnan()).all()); same_mask = bool(((a[6] == 0) == (b[6] == 0)).all())
print("fx vs strict: max rel err per component", ["%.1e" % e for e in err], "NaN pattern same", same_nan, "mask same", same_mask, "fixups", rb.fixups)
print(f"{os.environ.get('OLRT_LIB','default'):40s} n={n} strict {ms:7.3f} ms (+resid {msr:7.3f}) ({n*12/ms/1e6:6.1f} G rs/s)  fast(fx) {msf:7.3f}  fast(replay) {msx:7.3f}  fp32(fx) {ms32:7.3f} fp32(replay) {ms32r:7.3f}")
