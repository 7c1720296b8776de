"""TEST INFRASTRUCTURE ONLY -- the CPU baseline arm: the NumPy oracle (oracle/trace_np.py, the
restatement of the reference's NumPy float64 path) timed on the host cores.

The reference is single-threaded NumPy; "all the host threads it can use" is one worker process
per core, each tracing its own chunk of rays (rays are independent), exactly the courtesy
"all cores" figure of BASELINE.md section 3. One step = `chunks` x `chunk_rays` rays spread over the
workload's fields: ray launch (oracle/raygen_np.py) + trace + RMS-spot sums.

    python -m oracle.cpu_bench --lens even_asphere_double_gauss --steps 1 --budget 15
prints one JSON object. bench.py runs it as a subprocess so that no worker is forked from a
process that holds a CUDA context.
"""
from __future__ import annotations

import argparse
import json
import multiprocessing as mp
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import raygen_np, trace_np  # noqa: E402

_G = {}


def _setup(lens_name, wavelength):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from optiland_pr_b200 import samples
    from optiland_pr_b200.prescription import pack_surface_group

    lens = {"even_asphere_double_gauss": samples.EvenAsphereDoubleGauss,
            "cooke_triplet": samples.CookeTriplet, "double_gauss": samples.DoubleGauss}[lens_name]()
    packed = pack_surface_group(lens.surface_group.surfaces, [wavelength], record="last")
    surf_z = [float(s.geometry.cs.z) for s in lens.surface_group.surfaces[1:]]
    _G.update(packed=packed, surf_z=surf_z, EPL=lens._EPL, EPD=lens._EPD, max_field=lens.max_field,
              wavelength=wavelength, S_eff=lens.surface_group.num_surfaces - 1)


def _work(task):
    """One chunk: launch + trace + partial spot sums. Returns (count, Sx, Sy, Sxx, Syy, field)."""
    field_index, (hx, hy), n, seed = task
    px, py = raygen_np.random_pupil(n, seed)
    comps = raygen_np.generate_rays(hx, hy, px, py, _G["wavelength"], EPL=_G["EPL"], EPD=_G["EPD"],
                                    max_field=_G["max_field"], surface_z=_G["surf_z"])
    rays = trace_np.Rays(*comps)
    trace_np.trace_packed(_G["packed"], rays)
    x, y = rays.x, rays.y
    return (float(x.size), float(x.sum()), float(y.sum()), float((x * x).sum()), float((y * y).sum()),
            field_index)


def run(lens_name, fields, wavelength, steps=1, warmup=0, budget_s=15.0, chunk_rays=250_000, cores=None):
    cores = cores or len(os.sched_getaffinity(0))
    _setup(lens_name, wavelength)
    S_eff = _G["S_eff"]
    ctx = mp.get_context("fork")
    with ctx.Pool(cores) as pool:
        # calibrate: one chunk per core
        t0 = time.perf_counter()
        pool.map(_work, [(k % len(fields), fields[k % len(fields)], chunk_rays, 1000 + k) for k in range(cores)])
        t_cal = time.perf_counter() - t0
        rounds = max(1, int(budget_s / max(t_cal, 1e-3) / max(steps + warmup, 1)))
        rounds = min(rounds, 8)
        tasks = [(k % len(fields), fields[k % len(fields)], chunk_rays, k) for k in range(cores * rounds)]
        for _ in range(warmup):
            pool.map(_work, tasks)
        t0 = time.perf_counter()
        for _ in range(steps):
            parts = pool.map(_work, tasks)
        dt = (time.perf_counter() - t0) / steps
    rays_per_step = len(tasks) * chunk_rays
    rms = []
    for f in range(len(fields)):
        p = np.array([q[:5] for q in parts if q[5] == f]).sum(axis=0)
        if p.size and p[0] > 0:
            mx, my = p[1] / p[0], p[2] / p[0]
            rms.append(float(np.sqrt(max((p[3] + p[4]) / p[0] - mx * mx - my * my, 0.0))))
    return dict(value=rays_per_step * S_eff / dt, ms_per_step=dt * 1e3, cores=cores, rms=rms,
                sample=f"{rays_per_step} rays ({len(tasks)} chunks of {chunk_rays}) over {len(fields)} fields per step, "
                       f"NumPy fp64 oracle port, {cores} worker processes")


def run_subprocess(lens_name, fields, wavelength, steps=1, warmup=0, budget_s=15.0):
    import subprocess

    cmd = [sys.executable, "-m", "oracle.cpu_bench", "--lens", lens_name, "--steps", str(steps),
           "--warmup", str(warmup), "--budget", str(budget_s), "--wavelength", repr(wavelength),
           "--fields", json.dumps(fields)]
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", OMP_NUM_THREADS="1", OPENBLAS_NUM_THREADS="1",
               MKL_NUM_THREADS="1")
    out = subprocess.run(cmd, cwd=ROOT, env=env, check=True, capture_output=True, text=True).stdout
    return json.loads(out.strip().splitlines()[-1])


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--lens", default="even_asphere_double_gauss")
    ap.add_argument("--steps", type=int, default=1)
    ap.add_argument("--warmup", type=int, default=0)
    ap.add_argument("--budget", type=float, default=15.0)
    ap.add_argument("--wavelength", type=float, default=0.5876)
    ap.add_argument("--fields", default="[[0.0, 0.0], [0.0, 0.7142857142857143], [0.0, 1.0]]")
    a = ap.parse_args()
    print(json.dumps(run(a.lens, [tuple(f) for f in json.loads(a.fields)], a.wavelength, a.steps, a.warmup,
                         a.budget)))
