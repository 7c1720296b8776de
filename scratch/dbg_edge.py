import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
from test_gpu_edge import _singlet, INF
from oracle import trace_np
from optiland_pr_b200.prescription import pack_surface_group
from optiland_pr_b200.trace import fused_trace
lens = _singlet(radius=INF)
n = 32
N = np.concatenate([np.array([0.0, 1e-15, -1e-15, 1e-14, 2e-14, 1e-9]), np.linspace(0.01, 1.0, n - 6)])
L = np.sqrt(np.maximum(0.0, 1.0 - N * N))
rays_in = np.stack([np.zeros(n), np.linspace(-1, 1, n), np.full(n, -1.0), L, np.zeros(n), N, np.ones(n), np.full(n, 0.55)])
packed = pack_surface_group(lens.surface_group.surfaces, [0.55])
with np.errstate(all="ignore"):
    want = np.array(trace_np.trace_packed(packed, trace_np.Rays(*rays_in))["records"])
for fast in (False, True):
    res = fused_trace(packed, [torch.tensor(rays_in[q], device="cuda") for q in range(8)], fast=fast)
    got = res.record.cpu().numpy()
    bad = np.argwhere(np.isnan(got) != np.isnan(want))
    print("fast", fast, "fixups", res.fixups, "mismatch", bad.tolist()[:12])
    for s, q, r in bad[:6]:
        print("  surf", s, "comp", q, "ray", r, "got", got[s, :, r], "want", want[s, :, r])
