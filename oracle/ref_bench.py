"""TEST INFRASTRUCTURE ONLY -- the reference arm of bench.py: the UNMODIFIED reference timed on this box.

Imports the reference itself (oracle/ref_shim.py: /root/reference in the build container, the verbatim copy under
oracle/_ref on the GPU box) and times its own public API on the bench workload (BASELINE.json configs[1]: the
even-asphere double Gauss, 3 fields, primary wavelength, random pupil; `Optic.trace` per field + the RMS spot of
RayOperand.rms_spot_size, optimization/operand/ray.py:336-340):

  numpy   -- `be.set_backend("numpy")`, the reference's CPU path. `single`: one Python process, as shipped (NumPy
             elementwise ops are single-threaded). `all`: one worker process per host core, each tracing its own
             chunk of the same workload through the same unmodified code -- all the host threads the reference can use.
  torch   -- `be.set_backend("torch"); be.set_device("cuda")`: the reference's torch-eager path on the GPU (one CUDA
             kernel per elementwise op, a host sync per Newton iteration, newton_raphson.py:148 /
             torch_backend.py:451-454) -- the only thing README.md:120 ("150M+ ray-surfaces/s") can refer to.

    python -m oracle.ref_bench numpy --steps 1 --budget 20
    python -m oracle.ref_bench torch --precision float64 --lens c2 --rays 1000000
Each prints one JSON object. bench.py runs them as subprocesses (no worker is forked from a CUDA process).
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

from oracle import ref_shim  # noqa: E402

FIELDS = [(0.0, 0.0), (0.0, 10.0 / 14.0), (0.0, 1.0)]
WAVELENGTH = 0.5876

# BASELINE C2 lens (the stock double Gauss with surfaces 1, 5, 7, 11 as even aspheres); the same
# coefficients as optiland_pr_b200/samples.py: DOUBLE_GAUSS_ASPHERES (tests/test_oracle_golden.py compares the packs)
_ASPHERES = {
    1: [0.0, -2.0e-7, 1.5e-10],
    5: [0.0, 1.2e-6, -3.0e-9],
    7: [0.0, -1.5e-6, 2.5e-9],
    11: [0.0, 1.0e-7, -0.5e-10],
}


def build_c2_lens():
    """The even-asphere double Gauss built with the reference's own Optic.add_surface."""
    import optiland.backend as be
    from optiland import optic as ref_optic
    aspheres = _ASPHERES
    lens = ref_optic.Optic()
    rows = [
        dict(index=1, radius=56.20238, thickness=8.75, material="N-SSK2"),
        dict(index=2, radius=152.28580, thickness=0.5),
        dict(index=3, radius=37.68262, thickness=12.5, material="N-SK2"),
        dict(index=4, radius=be.inf, thickness=3.8, material=("F5", "schott")),
        dict(index=5, radius=24.23130, thickness=16.369445),
        dict(index=6, radius=be.inf, thickness=13.747957, is_stop=True),
        dict(index=7, radius=-28.37731, thickness=3.8, material=("F5", "schott")),
        dict(index=8, radius=be.inf, thickness=11, material="N-SK16"),
        dict(index=9, radius=-37.92546, thickness=0.5),
        dict(index=10, radius=177.41176, thickness=7, material="N-SK16"),
        dict(index=11, radius=-79.41143, thickness=61.487536),
    ]
    lens.add_surface(index=0, radius=be.inf, thickness=be.inf)
    for row in rows:
        if row["index"] in aspheres:
            row = dict(row, surface_type="even_asphere", conic=0.0, coefficients=list(aspheres[row["index"]]))
        lens.add_surface(**row)
    lens.add_surface(index=12)
    lens.set_aperture(aperture_type="imageFNO", value=5)
    lens.set_field_type(field_type="angle")
    for fy in (0, 10, 14):
        lens.add_field(y=fy)
    lens.add_wavelength(value=0.4861)
    lens.add_wavelength(value=0.5876, is_primary=True)
    lens.add_wavelength(value=0.6563)
    return lens


def build_lens(name):
    if name == "c2":
        return build_c2_lens(), FIELDS, [WAVELENGTH]
    if name == "c1":
        from optiland.samples.objectives import CookeTriplet
        lens = CookeTriplet()
        return lens, [(0.0, 0.0), (0.0, 0.7), (0.0, 1.0)], [0.48, 0.55, 0.65]
    raise KeyError(name)


_G = {}


def _trace_chunk(task):
    """One `Optic.trace` of the unmodified reference + the RMS-spot sums of its image-surface rows."""
    field_index, n, seed = task
    import optiland.backend as be
    from optiland.distribution import RandomDistribution
    lens = _G["lens"]
    hx, hy = _G["fields"][field_index]
    dist = RandomDistribution(seed=seed)
    dist.generate_points(n)
    with np.errstate(all="ignore"):
        lens.trace(hx, hy, _G["wl"], None, dist)
    sg = lens.surface_group
    x, y = be.to_numpy(sg.x[-1, :]), be.to_numpy(sg.y[-1, :])
    return (float(x.size), float(x.sum()), float(y.sum()), float((x * x).sum()), float((y * y).sum()), field_index)


def _rms(parts, nf):
    out = []
    for f in range(nf):
        p = np.array([q[:5] for q in parts if q[5] == f]).sum(axis=0)
        if np.size(p) and p[0] > 0:
            mx, my = p[1] / p[0], p[2] / p[0]
            out.append(float(np.sqrt(max((p[3] + p[4]) / p[0] - mx * mx - my * my, 0.0))))
    return out


def run_numpy(lens_name="c2", steps=1, warmup=0, budget_s=20.0, chunk_rays=250_000, cores=None, mode="all"):
    ref_shim.install()
    import optiland.backend as be
    be.set_backend("numpy")
    lens, fields, wls = build_lens(lens_name)
    _G.update(lens=lens, fields=fields, wl=wls[len(wls) // 2] if lens_name == "c1" else wls[0])
    S_eff = len(lens.surface_group.surfaces) - 1
    nf = len(fields)
    cores_avail = len(os.sched_getaffinity(0))
    cores = 1 if mode == "single" else (cores or cores_avail)
    if mode == "single":
        t0 = time.perf_counter()
        _trace_chunk((0, 50_000, 999))
        t_cal = (time.perf_counter() - t0) / 50_000
        per_field = int(max(100_000, min(2_000_000, budget_s / max(steps + warmup, 1) / nf / t_cal)))
        per_field = min(per_field, 400_000) if budget_s <= 25 else per_field
        tasks = [(f, per_field, 1 + f) for f in range(nf)]
        for _ in range(warmup):
            [_trace_chunk(t) for t in tasks]
        t0 = time.perf_counter()
        for _ in range(steps):
            parts = [_trace_chunk(t) for t in tasks]
        dt = (time.perf_counter() - t0) / steps
        rays_per_step = per_field * nf
        sample = (f"{rays_per_step} rays ({nf} fields x {per_field}) per step, UNMODIFIED reference, NumPy backend, "
                  f"one Python process as shipped ({cores_avail} host cores available)")
    else:
        ctx = mp.get_context("fork")
        with ctx.Pool(cores) as pool:
            t0 = time.perf_counter()
            pool.map(_trace_chunk, [(k % nf, chunk_rays, 1000 + k) for k in range(cores)])
            t_cal = time.perf_counter() - t0
            rounds = max(1, min(8, int(budget_s / max(t_cal, 1e-3) / max(steps + warmup, 1))))
            tasks = [(k % nf, chunk_rays, k) for k in range(cores * rounds)]
            for _ in range(warmup):
                pool.map(_trace_chunk, tasks)
            t0 = time.perf_counter()
            for _ in range(steps):
                parts = pool.map(_trace_chunk, tasks)
            dt = (time.perf_counter() - t0) / steps
        rays_per_step = len(tasks) * chunk_rays
        sample = (f"{rays_per_step} rays ({len(tasks)} chunks of {chunk_rays}) over {nf} fields per step, UNMODIFIED "
                  f"reference, NumPy backend, {cores} worker processes (one Optic.trace per chunk)")
    return dict(value=rays_per_step * S_eff / dt, ms_per_step=dt * 1e3, cores=cores, rms=_rms(parts, nf),
                sample=sample, rays_per_step=rays_per_step, kind="reference")


def run_torch(lens_name="c2", precision="float64", rays=1_000_000, steps=3, warmup=1):
    """The reference's torch-eager path on the GPU: one Optic.trace per (field, wavelength) pair + RMS spot."""
    import torch
    ref_shim.install()
    import optiland.backend as be
    be.set_backend("torch")
    be.set_device("cuda")
    be.set_precision(precision)
    torch.set_default_device("cuda")   # the reference needs it: be.atleast_1d ignores the configured device
    lens, fields, wls = build_lens(lens_name)
    from optiland.distribution import RandomDistribution
    S_eff = len(lens.surface_group.surfaces) - 1
    dist = RandomDistribution(seed=1)
    dist.generate_points(rays)
    pairs = [(f, w) for f in fields for w in wls]

    def step():
        out = []
        for (hx, hy), w in pairs:
            lens.trace(hx, hy, w, None, dist)
            sg = lens.surface_group
            x, y = sg.x[-1, :], sg.y[-1, :]
            out.append(torch.sqrt(torch.mean((x - x.mean()) ** 2 + (y - y.mean()) ** 2)))
        return [float(v) for v in out]

    if warmup:     # context creation, lazy kernel loading, allocator growth: one pass on a tenth of the rays
        small = RandomDistribution(seed=2)
        small.generate_points(max(1000, rays // 10))
        dist, small = small, dist
        step()
        dist, small = small, dist
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        rms = step()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / steps
    return dict(value=rays * len(pairs) * S_eff / dt, ms_per_step=dt * 1e3, rms=rms, precision=precision,
                lens=lens_name, rays_per_pair=rays, pairs=len(pairs),
                sample=f"{len(pairs)} x Optic.trace of {rays} rays, UNMODIFIED reference, torch backend on cuda "
                       f"({precision}), RMS spot -> host")


def run_subprocess(argv, timeout=900):
    import subprocess
    env = dict(os.environ, OMP_NUM_THREADS="1", OPENBLAS_NUM_THREADS="1", MKL_NUM_THREADS="1")
    if argv[0] == "numpy":
        env["CUDA_VISIBLE_DEVICES"] = ""
    p = subprocess.run([sys.executable, "-m", "oracle.ref_bench"] + [str(a) for a in argv], cwd=ROOT, env=env,
                       capture_output=True, text=True, timeout=timeout)
    if p.returncode != 0:
        raise RuntimeError("oracle.ref_bench failed: " + p.stderr[-600:])
    return json.loads(p.stdout.strip().splitlines()[-1])


def available() -> bool:
    return ref_shim.reference_available()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("what", choices=["numpy", "torch"])
    ap.add_argument("--lens", default="c2")
    ap.add_argument("--mode", default="all", choices=["all", "single"])
    ap.add_argument("--steps", type=int, default=1)
    ap.add_argument("--warmup", type=int, default=0)
    ap.add_argument("--budget", type=float, default=20.0)
    ap.add_argument("--precision", default="float64")
    ap.add_argument("--rays", type=int, default=1_000_000)
    a = ap.parse_args()
    import warnings
    warnings.simplefilter("ignore")
    if a.what == "numpy":
        r = run_numpy(a.lens, a.steps, a.warmup, a.budget, mode=a.mode)
    else:
        r = run_torch(a.lens, a.precision, a.rays, a.steps, a.warmup)
    print(json.dumps(r))
