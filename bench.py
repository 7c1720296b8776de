#!/usr/bin/env python
"""bench.py -- ray-surface intersections per second of the fused real-ray trace on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W]            # this repo's CUDA path
    python bench.py --impl reference [--steps K] [--warmup W]      # CPU baseline arm (NumPy port)
    torchrun --nproc-per-node N bench.py --gpus N ...              # one rank per GPU

Workload (BASELINE.json configs[1]): the even-asphere double Gauss (S_eff = 12 intersected
surfaces, 4 of them Newton-Raphson even aspheres), 3 fields x 16 777 216 random pupil rays at the
primary wavelength, fp64, forward trace + RMS-spot epilogue. One step = one pass over that batch.
`value` = rays x S_eff / time with the pupil grid resident in HBM; `e2e` = the same through the public
call in the reference's own shape, `Optic.trace(Hx, Hy, wl, num_rays, "random", seed)`: the named distribution
is drawn on the device from its seed every step (NumPy's PCG64 stream), traced, reduced and the RMS values read back
to the host -- all inside the timed region; `e2e.host_grid` is the same with the caller's own pupil grid in pinned
HOST memory (`distribution=(Px, Py)`: 16 bytes per pupil point cross PCIe every step, streamed under the kernel).
Multi-GPU: rays are independent, every rank traces its own batch
(weak scaling); the only collective is an NCCL all-reduce of the [3][10] spot sums.

Only the `cpu_baseline` / `extras.reference_torch_eager_cuda` legs and `--impl reference` execute oracle/: they run
the UNMODIFIED reference (oracle/ref_bench.py; on the GPU box the verbatim copy under oracle/_ref that
__graft_entry__.build() makes) -- NumPy backend on the host cores, and its torch-eager path on the same GPU -- and the
NumPy port (oracle/cpu_bench.py) as a second, faster CPU figure.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

RAYS_PER_FIELD = 16_777_216
FIELDS = [(0.0, 0.0), (0.0, 10.0 / 14.0), (0.0, 1.0)]
WAVELENGTH = 0.5876
METRIC = "ray-surfaces/sec fwd (fp64, even-asphere double Gauss, 3 fields x 16M rays) on B200"
UNIT = "ray-surfaces/s"
# the SAME string in both arms (the reference arm states its bounded sub-sample under config.sample only)
WORKLOAD = ("even-asphere double Gauss (12 intersected surfaces, 4 Newton-Raphson even aspheres), 3 fields x 16777216 "
            "random pupil rays per GPU, primary wavelength, ray launch + fwd trace + final ray state + RMS spot")


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--rays-per-field", type=int, default=RAYS_PER_FIELD)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--extras", type=int, default=1, help="also time fp32 / fast / fwd+bwd legs")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------
# clocks: nvidia-smi sampled DURING the timed region
# ---------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.samples, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            parts = [p.strip() for p in s.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0]))
                mx = float(parts[1])
            except ValueError:
                continue
            for nm, v in zip(names, parts[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------
# CPU arm: the NumPy oracle on a bounded sample of the same workload, all host cores
# ---------------------------------------------------------------------------------------------
def cpu_arm(steps: int, warmup: int, budget_s: float = 20.0):
    from oracle import cpu_bench

    return cpu_bench.run_subprocess("even_asphere_double_gauss", [list(f) for f in FIELDS], WAVELENGTH,
                                    steps=steps, warmup=warmup, budget_s=budget_s)


def reference_cpu(steps: int, warmup: int, budget_all: float, budget_single: float):
    """The UNMODIFIED reference on the host cores (oracle/ref_bench.py): every core (one worker process per core, each
    tracing its own chunk through the reference's Optic.trace) and one process as shipped."""
    from oracle import ref_bench
    if not ref_bench.available():
        return None
    r_all = ref_bench.run_subprocess(["numpy", "--mode", "all", "--steps", steps, "--warmup", warmup,
                                      "--budget", budget_all])
    r_one = ref_bench.run_subprocess(["numpy", "--mode", "single", "--steps", 1, "--budget", budget_single])
    return {"value": r_all["value"], "unit": UNIT, "cores": r_all["cores"], "kind": "reference",
            "sample": r_all["sample"], "ms_per_step": r_all["ms_per_step"], "rms_spot_mm": r_all["rms"],
            "single_process": {"value": r_one["value"], "unit": UNIT, "cores": 1, "sample": r_one["sample"],
                               "rms_spot_mm": r_one["rms"]}}


def reference_main(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps = max(1, min(args.steps, 3))
    warm = min(args.warmup, 1)
    r = reference_cpu(steps, warm, budget_all=40.0, budget_single=15.0)
    if r is None:   # no reference on this box (oracle/_ref not built): the NumPy port of it
        p = cpu_arm(steps, warm, budget_s=30.0)
        r = {"value": p["value"], "unit": UNIT, "cores": p["cores"], "kind": "port", "sample": p["sample"],
             "ms_per_step": p["ms_per_step"]}
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": warm, "ms_per_step": r["ms_per_step"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": r["sample"]},
        "cpu_baseline": {k: v for k, v in r.items() if k != "ms_per_step"},
        "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# ---------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------
def main():
    args = parse()
    if args.impl == "reference":
        reference_main(args)
        return
    import torch
    import torch.distributed as dist

    import optiland_pr_b200 as ol
    from optiland_pr_b200 import _lib, flops, parallel, samples, spot
    from optiland_pr_b200.trace import fused_trace

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa_cpus = parallel.bind_to_gpu_numa(local) if world > 1 else None   # before any pinned allocation
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.load()
    ol.backend.set_device(f"cuda:{local}")
    ol.backend.set_precision("float64")

    npf = args.rays_per_field
    nf = len(FIELDS)
    lens = samples.EvenAsphereDoubleGauss()
    lens.surface_group.record = "none"
    S_eff = lens.surface_group.num_surfaces - 1
    # pupil grid on the host (pinned): the reference's RandomDistribution(seed) arithmetic
    rng = np.random.default_rng(1 + rank)
    r_ = rng.uniform(size=npf)
    th = rng.uniform(0, 2 * np.pi, size=npf)
    Px_h = torch.from_numpy(np.sqrt(r_) * np.cos(th)).pin_memory()
    Py_h = torch.from_numpy(np.sqrt(r_) * np.sin(th)).pin_memory()
    del r_, th
    Hx = [f[0] for f in FIELDS]
    Hy = [f[1] for f in FIELDS]

    # ---- device-resident inputs for `value`: the pupil grid (what Optic.trace is given), already in HBM
    from optiland_pr_b200.trace import fused_trace_pupil, group_true_iters
    Px_d, Py_d = Px_h.to(dev), Py_h.to(dev)
    n = nf * npf
    packed = lens.surface_group.pack([WAVELENGTH])
    fxy_np, z_start = lens._field_offsets(np.asarray(Hx), np.asarray(Hy))
    fxy_d = torch.from_numpy(fxy_np).to(dev)
    gen_kw = dict(EPD=lens._EPD, EPL=lens._EPL, z_start=z_start, wavelength=WAVELENGTH, dtype=torch.float64)
    out = torch.empty((8, n), dtype=torch.float64, device=dev)
    sums = torch.zeros((nf, _lib.SPOT_SUMS), dtype=torch.float64, device=dev)
    status_t = torch.zeros(1, dtype=torch.int32, device=dev)
    # Newton trip counts PER FIELD (the reference's Optic.trace takes one field per call, so each field's rays decide
    # their own exit test, newton_raphson.py:148): kf [nf][S] goes to the kernel as `group_k`, the per-field reports
    # come back in `resid` [nf][S][GROUP_RESID]
    kf = np.tile(np.asarray(packed.nr_iters(), dtype=np.int32), (nf, 1))
    group_k = [torch.from_numpy(kf.copy()).to(dev)]
    resid = torch.zeros((nf, packed.n_surfaces, _lib.GROUP_RESID), dtype=torch.int64, device=dev)
    zero1 = torch.zeros(1, dtype=torch.float64, device=dev)
    chief = torch.empty((8, nf), dtype=torch.float64, device=dev)

    # multi-GPU: the merit exchange of step k (one all-gather of the [3][10] sums, < 1 KB) runs on NCCL's stream
    # while step k+1 traces; it is merged one step later (double-buffered)
    sums2 = [torch.zeros_like(sums) for _ in range(2)]
    parts2 = [[torch.empty_like(sums) for _ in range(world)] for _ in range(2)]
    works = [None, None]

    def merge(slot_):
        works[slot_].wait()
        works[slot_] = None
        buf = torch.stack(parts2[slot_])
        m = buf.sum(dim=0)
        m[:, 5] = buf[:, :, 5].amax(dim=0)
        m[:, 7:] = sums2[slot_][:, 7:]
        sums.copy_(m)

    resid_h = [torch.zeros(resid.shape, dtype=torch.int64).pin_memory() for _ in range(2)]
    resid_ev = [None, None]
    tick = [0]

    def check_newton(slot):
        """The reference's exit rule (newton_raphson.py:148) on the residual maxima one step reported."""
        resid_ev[slot].synchronize()
        return bool((group_true_iters(packed, resid_h[slot].numpy().view(np.float64), kf) == kf).all())

    shift_box = [None]

    def step_resident(settle=False):
        """One pass of the hot path: launch the rays of the 3 fields from the pupil grid, trace them, write the
        final ray state and take the spot sums -- one kernel (olrt_trace_pupil_f64) plus its IEEE fix-up pass.
        The Newton trip counts every step used are checked against the reference's exit rule from the residual
        maxima the kernel reports; the (small) read-back of step k is examined while step k+1 runs."""
        slot = tick[0] & 1
        cur = sums2[slot] if world > 1 else sums
        cur.zero_()
        resid.zero_()
        if settle or shift_box[0] is None:
            # the point the spot sums are taken about (conditioning of the second moments only: the statistics do
            # not depend on it): each field's chief-ray image point, traced once per lens, not once per step
            fused_trace_pupil(packed, zero1, zero1, fxy_d, out=chief, **gen_kw)
            shift_box[0] = chief[:2].t().nan_to_num(0.0, 0.0, 0.0).contiguous()
        shift = shift_box[0]
        fused_trace_pupil(packed, Px_d, Py_d, fxy_d, out=out, spot=cur, spot_shift=shift, group_k=group_k[0],
                          group_resid=resid, status_t=status_t, **gen_kw)
        cur[:, 8:10] = shift
        if world > 1:
            works[slot] = dist.all_gather(parts2[slot], cur, async_op=True)
            if works[slot ^ 1] is not None:
                merge(slot ^ 1)
        resid_h[slot].copy_(resid, non_blocking=True)
        resid_ev[slot] = torch.cuda.Event()
        resid_ev[slot].record()
        prev = slot ^ 1
        if resid_ev[prev] is not None and not settle:
            if not check_newton(prev):
                raise SystemExit("Newton trip counts changed between steps")
            resid_ev[prev] = None
        tick[0] += 1
        if settle:   # first calls: find the counts the exit rule asks for
            resid_ev[slot].synchronize()
            kt = group_true_iters(packed, resid_h[slot].numpy().view(np.float64), kf)
            resid_ev[slot] = None
            if not (kt == kf).all():
                kf[:] = np.where(kt < 0, np.minimum(kf * 2 + 1, _lib.GROUP_RESID - 1), kt)
                packed.set_nr_iters(kf.max(axis=0).tolist())   # the kernel's slot layout covers every field's counts
                group_k[0] = torch.from_numpy(kf.copy()).to(dev)
                return False
        return True

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms / steps

    # settle the Newton trip counts K_s (first call may relaunch), then time
    for _ in range(8):
        if step_resident(settle=True):
            break
    nr_iters = packed.nr_iters()
    nr_iters_per_field = kf.tolist()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms_step = timed(step_resident, args.steps, max(args.warmup, 3))
    clocks = sampler.stop() if rank == 0 else None
    for slot in (0, 1):   # the last step's residuals and merit exchange
        if resid_ev[slot] is not None and not check_newton(slot):
            raise SystemExit("Newton trip counts changed between steps")
    last = (tick[0] - 1) & 1
    for slot in (last ^ 1, last):
        if works[slot] is not None:
            merge(slot)
    launches_per_step = 2  # the fused launch / trace / spot kernel + its IEEE fix-up pass (strict build)
    value = n * S_eff * world / (ms_step * 1e-3)
    rms_resident = spot.spot_stats(sums)["rms"].cpu().tolist()
    if int(status_t.item()) != 0:
        raise SystemExit("trace status word is not clean")

    # ---- parity guard on the bench batch itself: the timed (tolerance-mode, in-kernel launched) result against
    # the separate launch kernel + the strict, bit-faithful trace kernel (the pair tests/ pin to the reference bit
    # for bit) on the same rays
    rays = lens.launch_rays(np.asarray(Hx), np.asarray(Hy), Px_d, Py_d, WAVELENGTH)
    ins = rays.components()
    ins[7] = None
    # field by field, as the reference's Optic.trace is called: each field's rays settle their own Newton trip counts
    # (newton_raphson.py:148); the one-launch step must have used exactly those (`nr_iters_coincide`)
    strict_out = torch.empty_like(out)
    pk_f = lens.surface_group.pack([WAVELENGTH])
    per_field = []
    for f_ in range(nf):
        sl = slice(f_ * npf, (f_ + 1) * npf)
        res_f = fused_trace(pk_f, [c[sl] if c is not None else None for c in ins], out=strict_out[:, sl], fast=False)
        per_field.append([int(k) for k in res_f.nr_iters])
    den = torch.clamp(strict_out.abs(), min=1.0)
    parity = {"max_rel_err_vs_strict_kernel": float(((out - strict_out).abs() / den).nan_to_num(0.0).max()),
              "nan_pattern_identical": bool((out.isnan() == strict_out.isnan()).all()),
              "vignetting_mask_identical": bool(((out[6] == 0) == (strict_out[6] == 0)).all()),
              "bar": "1e-10 relative, masks identical (BASELINE.json north_star)"}
    del den
    parity["nr_iters_per_field"] = nr_iters_per_field
    parity["nr_iters_field_by_field_calls"] = per_field
    parity["nr_iters_coincide"] = per_field == nr_iters_per_field
    if not parity["nr_iters_coincide"]:
        raise SystemExit(f"Newton trip counts of the one-launch step {nr_iters_per_field} differ from those of "
                         f"field-by-field calls {per_field}")
    # ---- the same check where the masks are not trivial ("C2 with aperture= radial stops"): the lens with
    # radial stops on its first, stop and last glass surfaces vignettes part of every off-axis bundle; tolerance-mode
    # kernel vs the strict kernel on the same 50 M rays: every clip decision must be the same one
    from optiland_pr_b200.apertures import RadialAperture
    lens_s = samples.EvenAsphereDoubleGauss()
    lens_s.surface_group.record = "none"
    sfs_s = lens_s.surface_group.surfaces
    sfs_s[1].aperture = RadialAperture(r_max=20.5)
    sfs_s[6].aperture = RadialAperture(r_max=6.2, r_min=0.05)
    sfs_s[11].aperture = RadialAperture(r_max=16.5)
    packed_s = lens_s.surface_group.pack([WAVELENGTH])
    packed_s.set_nr_iters(packed.nr_iters())
    out_s = fused_trace(packed_s, ins, fast=True).out
    fused_trace(packed_s, ins, out=strict_out, fast=False)
    clipped = (strict_out[6] == 0)
    den = torch.clamp(strict_out.abs(), min=1.0)
    parity["stopped_lens"] = {
        "radial_stops": "r_max 20.5 (surface 1), 0.05 <= r <= 6.2 (stop), r_max 16.5 (surface 11)",
        "rays": int(n), "clipped_rays": int(clipped.sum()), "clipped_fraction": float(clipped.double().mean()),
        "vignetting_mask_identical": bool(((out_s[6] == 0) == clipped).all()),
        "nan_pattern_identical": bool((out_s.isnan() == strict_out.isnan()).all()),
        "max_rel_err_vs_strict_kernel": float(((out_s - strict_out).abs() / den).nan_to_num(0.0).max())}
    del den, out_s, clipped
    ms_strict = timed(lambda: fused_trace(packed, ins, out=strict_out, fast=False, verify_nr=False, check_status=False),
                      max(2, args.steps // 2), 3)
    ms_fx_arrays = timed(lambda: fused_trace(packed, ins, out=strict_out, fast=True, verify_nr=False, check_status=False),
                         max(2, args.steps // 2), 3)
    del strict_out, rays, ins

    # ---- dominant kernel alone (roofline): trace_fx_kernel<double> in pupil mode with the fused spot sums
    shift0 = sums[:, 8:10].contiguous()

    def only_trace():
        fused_trace_pupil(packed, Px_d, Py_d, fxy_d, out=out, spot=sums, spot_shift=shift0, group_k=group_k[0],
                          group_resid=resid, **gen_kw)
    leg_clock = ClockSampler(local)      # the SM clock DURING this leg: the arithmetic peak below is quoted at it
    leg_clock.start()
    ms_kernel = timed(only_trace, max(args.steps, 40), 3)
    leg_clocks = leg_clock.stop()

    def only_trace_unverified():   # the same launch without the Newton residual reports (counts trusted, not checked)
        fused_trace_pupil(packed, Px_d, Py_d, fxy_d, out=out, spot=sums, spot_shift=shift0, **gen_kw)
    ms_unverified = timed(only_trace_unverified, args.steps, 3)
    fl_ray = float(np.mean([flops.flops_per_ray(packed, k) for k in nr_iters_per_field]))   # equal ray counts per field
    by_ray = 64.0 + 16.0 / nf   # 8 output doubles per ray + the pupil point (Px, Py) shared by the nf fields
    # measured FP64 FMA peak on this GPU (dependent-free DFMA chains, all SMs)
    sink = torch.zeros(1, dtype=torch.float64, device=dev)
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    st = torch.cuda.current_stream().cuda_stream

    def peak(fn, iters, blocks, threads):
        def go():
            _lib.check(fn(sink.data_ptr() if fn is lib.olrt_fma_peak_f64 else sink32.data_ptr(), iters, blocks,
                          threads, st), "fma_peak")
        ms = timed(go, 3, 2)
        return blocks * threads * iters * 8 * 2 / (ms * 1e-3) / 1e12

    sink32 = torch.zeros(1, dtype=torch.float32, device=dev)
    peak_clock = ClockSampler(local)
    peak_clock.start()
    fp64_peak = peak(lib.olrt_fma_peak_f64, 1 << 16, sms * 8, 256)
    for _ in range(6):                   # long enough for a few clock samples
        peak(lib.olrt_fma_peak_f64, 1 << 16, sms * 8, 256)
    peak_clocks = peak_clock.stop()
    fp32_peak = peak(lib.olrt_fma_peak_f32, 1 << 17, sms * 8, 256)
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    hbm_peak, hbm_src = 6650.0, "fallback"
    if os.path.exists(peaks_path):
        with open(peaks_path) as f:
            hbm_peak, hbm_src = float(json.load(f)["hbm_gbs"]), "measured"
    ach_tf = fl_ray * n / (ms_kernel * 1e-3) / 1e12
    ach_gbs = by_ray * n / (ms_kernel * 1e-3) / 1e9
    roofline = {
        "kernel": "trace_fx_kernel<double, ordinary-lens variant with per-field Newton counts and residual reports>, "
                  "pupil mode (in-kernel ray launch + trace + spot sums; tolerance mode: <= 1e-10 relative, masks "
                  "identical) + its IEEE fix-up pass: the launch of the timed step",
        "without_newton_reports": {
            "ms_per_launch": ms_unverified,
            "frac": flops.flops_per_ray(packed, nr_iters) * n / (ms_unverified * 1e-3) / 1e12 / fp64_peak,
            "flops_per_ray": flops.flops_per_ray(packed, nr_iters),
            "note": "the variant round 1's roofline was quoted on: batch-global trip counts taken on trust, no "
                    "evaluation of the exit test (the reference evaluates it every iteration)"},
        "bound": "fp64_fma", "achieved": ach_tf, "peak": fp64_peak, "unit": "TFLOP/s",
        "frac": ach_tf / fp64_peak, "peak_source": "olrt_fma_peak_f64 measured in this run",
        # 64 FP64 FMA lanes per SM and clock (B200_PROFILING.md): the arithmetic ceiling at the clock each leg ran at
        "peak_arithmetic": {"formula": "SMs x 64 DFMA/clk x 2 flop x SM clock",
                            "sms": sms, "sm_mhz_this_leg": leg_clocks.get("sm_mhz"),
                            "sm_mhz_peak_leg": peak_clocks.get("sm_mhz"),
                            "tflops_at_leg_clock": (sms * 64 * 2 * leg_clocks["sm_mhz"] * 1e6 / 1e12
                                                    if leg_clocks.get("sm_mhz") else None),
                            "tflops_at_peak_leg_clock": (sms * 64 * 2 * peak_clocks["sm_mhz"] * 1e6 / 1e12
                                                         if peak_clocks.get("sm_mhz") else None)},
        "flops_per_ray": fl_ray, "flops_note": "the trace's algorithmic count (SURVEY.md 8d); ray launch and spot "
                                               "sums are executed but not counted",
        "bytes_per_ray": by_ray, "nr_iters": nr_iters, "nr_iters_per_field": nr_iters_per_field, "ms_per_launch": ms_kernel,
        "hbm": {"bound": "hbm", "achieved": ach_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": ach_gbs / hbm_peak,
                "peak_source": hbm_src + " (MEASURED_PEAKS.json hbm_gbs)"},
        "traffic": None,
        "same_kernel_on_ray_arrays": {"ms_per_launch": ms_fx_arrays, "bytes_per_ray": 120.0,
                                      "note": "olrt_trace_fwd_f64 | OLRT_FWD_FAST: 7 input + 8 output arrays, no "
                                              "launch / spot fusion"},
        "strict_mode": {"kernel": "trace_fk_kernel<double> (-fmad=false, the reference's operation order, bit-identical)",
                        "ms_per_launch": ms_strict, "ray_surfaces_per_s": n * S_eff / (ms_strict * 1e-3),
                        "frac": fl_ray * n / (ms_strict * 1e-3) / 1e12 / fp64_peak},
    }
    tr_path = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tr_path):
        with open(tr_path) as f:
            roofline["traffic"] = json.load(f).get("trace_pupil_f64_bytes_per_launch")

    # ---- e2e: public API with host inputs
    h2d = Px_h.numel() * 8 * 2 + nf * 8 * 2
    # read back per step: the spot sums [nf][10] doubles, the status word, the Newton residual maxima
    d2h = nf * _lib.SPOT_SUMS * 8 + 4 + packed.n_surfaces * _lib.RESID_STRIDE * 8
    rms_host = torch.empty(nf, dtype=torch.float64).pin_memory()

    def finish_e2e(rr):
        if world > 1:
            # merit exchange (one < 1 KB all-reduce), then ONE read-back of the merged sums; the statistics on the host
            rms_host.copy_(spot.spot_stats(parallel.allreduce_spot_sums(rr.spot_sums).cpu())["rms"])
        else:   # the sums came back with the status word (one pinned read-back inside Optic.trace)
            rms_host.copy_(spot.spot_stats(rr.spot_sums_host)["rms"])

    def step_e2e():
        # the reference's own call shape, Optic.trace(Hx, Hy, wavelength, num_rays, "random") (optic/optic.py:584-609 ->
        # distribution.py:132-158): the pupil is a NAMED distribution, so what travels to the GPU is its seed -- the
        # generator's 128-bit state -- and the points are drawn on the device (olrt_pupil_random_f64), every step
        finish_e2e(lens.trace(Hx, Hy, WAVELENGTH, npf, "random", seed=1 + rank, spot=True))

    def step_e2e_grid():
        # the caller owns the pupil grid in (pinned) host memory: 16 bytes per pupil point cross PCIe every step
        finish_e2e(lens.trace(Hx, Hy, WAVELENGTH, distribution=(Px_h, Py_h), spot=True))

    del out, Px_d, Py_d
    torch.cuda.empty_cache()
    ms_e2e = timed(step_e2e, max(2, args.steps // 2), 3)
    e2e_value = n * S_eff * world / (ms_e2e * 1e-3)
    rms_e2e_named = rms_host.tolist()
    ms_e2e_grid = timed(step_e2e_grid, max(2, args.steps // 2), 3)
    h2d_named = 32 + nf * 8 * 2     # generator state + increment, the field table

    extras = {}
    if args.extras:
        extras["scaling_legs"] = scaling_legs(torch, dist, lens, packed, npf, nf, S_eff, timed, args, rank, world, dev)
    if args.extras and world == 1:
        extras.update(run_extras(torch, lens, packed, npf, nf, S_eff, timed, args, fp32_peak, fp64_peak, hbm_peak=hbm_peak))

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD if npf == RAYS_PER_FIELD else WORKLOAD.replace(str(RAYS_PER_FIELD), str(npf)),
                   "rays_per_gpu": n, "surfaces": S_eff, "precision": "float64, tolerance mode (<= 1e-10 relative to the reference, identical masks; "
                                                                  "the bit-identical strict mode is timed in roofline.strict_mode)",
                   "l2": "inputs (268 MB pupil grid) and outputs (3.2 GB) per step exceed the 126 MB L2; no flush needed",
                   "parallelism": f"rays sharded over {world} GPU(s), spot sums all-reduced",
                   "host_affinity": (f"each rank pinned to its GPU's NUMA CPUs ({len(numa_cpus)} on rank 0)"
                                     if numa_cpus else "default")},
        "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": ms_e2e, "h2d_bytes_per_step": h2d_named,
                "d2h_bytes_per_step": d2h,
                "api": "Optic.trace(Hx, Hy, wl, num_rays, 'random', seed, spot=True) + spot.spot_stats -> host: the "
                       "reference's call shape; the named distribution is drawn on the device from the seed (NumPy's "
                       "PCG64 stream, bit for bit) every step, traced, reduced, and the per-field RMS read back",
                "host_grid": {"value": n * S_eff * world / (ms_e2e_grid * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e_grid,
                              "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                              "api": "Optic.trace(Hx, Hy, wl, distribution=(Px, Py) pinned host, spot=True): the caller's "
                                     "own pupil grid; the copy is streamed in chunks on a side stream and overlaps the "
                                     "fused launch + trace + spot kernel of the chunks already on the device"}},
        "gpu_launches": launches_per_step * args.steps,
        "roofline": roofline, "parity_check": parity, "clocks": clocks,
        "rms_spot_mm": rms_resident, "rms_spot_mm_e2e": rms_e2e_named, "rms_spot_mm_e2e_host_grid": rms_host.tolist(),
        "extras": extras,
    }
    if not args.no_cpu_baseline and world == 1:
        r = cpu_arm(1, 0, budget_s=12.0)
        port = {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": r["sample"],
                "rms_spot_mm": r.get("rms")}
        ref = None
        try:
            ref = reference_cpu(1, 0, budget_all=15.0, budget_single=12.0)
        except Exception as e:   # a baseline leg must never take the bench line down
            port["reference_error"] = repr(e)[:300]
        if ref is not None:
            line["cpu_baseline"] = ref
            line["cpu_baseline"]["port_all_cores"] = port     # the NumPy restatement (oracle/trace_np.py), ~6x faster
        else:
            line["cpu_baseline"] = port
        if args.extras:
            line["extras"]["reference_torch_eager_cuda"] = reference_torch_eager()
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def scaling_legs(torch, dist, lens, packed, npf, nf, S_eff, timed, args, rank, world, dev):
    """The configurations BASELINE.json names for several GPUs, run at EVERY N (N = 1 is their baseline), so that the
    driver's per-N runs give their scaling curves:
      strong_c2  the C2 batch at FIXED total size (3 fields x 16.7 M rays) split over the ranks by pupil range, spot
                 sums all-reduced (the headline `value` is weak scaling: every rank its own 50 M rays);
      c5         Monte-Carlo tolerancing, 4096 trials x 262 144 rays, trials dealt to the ranks, merits gathered
                 (tolerancing/monte_carlo.py:60-123);
      c4         one Adam step of the RMS-spot merit on the Cooke triplet, 3 fields x 3 wavelengths x 1 M rays, the
                 (field, wavelength) pairs dealt round-robin to the ranks, parameter gradients all-reduced, the same
                 optimiser step on every rank (optimization/optimizer/torch/base.py:95-154)."""
    from optiland_pr_b200 import _lib, parallel, samples, spot
    from optiland_pr_b200.autograd import rms_merit_sharded
    from optiland_pr_b200.optic import random_pupil_device
    from optiland_pr_b200.trace import fused_trace_pupil
    out = {}
    # ---- strong scaling of C2
    Px, Py = random_pupil_device(npf, 1, dev)             # the same grid on every rank (same seed)
    lo, hi = parallel.shard_range(npf)
    Hx = np.asarray([f[0] for f in FIELDS]); Hy = np.asarray([f[1] for f in FIELDS])
    fxy_np, z_start = lens._field_offsets(Hx, Hy)
    fxy_d = torch.from_numpy(fxy_np).to(dev)
    gen_kw = dict(EPD=lens._EPD, EPL=lens._EPL, z_start=z_start, wavelength=WAVELENGTH, dtype=torch.float64)
    o = torch.empty((8, nf * npf), dtype=torch.float64, device=dev)
    sums = torch.zeros((nf, _lib.SPOT_SUMS), dtype=torch.float64, device=dev)
    zero1 = torch.zeros(1, dtype=torch.float64, device=dev)
    chief = torch.empty((8, nf), dtype=torch.float64, device=dev)
    fused_trace_pupil(packed, zero1, zero1, fxy_d, out=chief, **gen_kw)
    shift = chief[:2].t().nan_to_num(0.0, 0.0, 0.0).contiguous()
    resid = torch.zeros((packed.n_surfaces, _lib.RESID_STRIDE), dtype=torch.int64, device=dev)
    merged = [None]

    def strong():
        sums.zero_()
        resid.zero_()
        fused_trace_pupil(packed, Px, Py, fxy_d, p_lo=lo, p_len=hi - lo, out=o, spot=sums, spot_shift=shift,
                          resid_out=resid, **gen_kw)
        merged[0] = parallel.allreduce_spot_sums(sums)

    ms = timed(strong, args.steps, 3)
    out["strong_c2"] = {"ms_per_step": ms, "ray_surfaces_per_s": nf * npf * S_eff / (ms * 1e-3),
                        "rays_total": nf * npf, "rays_per_rank": nf * (hi - lo),
                        "rms_spot_mm": spot.spot_stats(merged[0])["rms"].cpu().tolist()}
    del o, Px, Py
    torch.cuda.empty_cache()
    # ---- C5
    out.update(bench_monte_carlo(torch, lens, S_eff))
    # ---- C4
    lens4 = samples.CookeTriplet()
    lens4.surface_group.record = "none"
    sfs = lens4.surface_group.surfaces
    pars = []
    for k in range(1, 7):
        p_ = torch.nn.Parameter(torch.tensor(float(sfs[k].geometry.radius), dtype=torch.float64, device=dev))
        lens4.set_radius(p_, k)
        pars.append(p_)
    for k in range(1, 7):
        p_ = torch.nn.Parameter(torch.tensor(0.0, dtype=torch.float64, device=dev))
        lens4.set_conic(p_, k)
        pars.append(p_)
    opt = torch.optim.Adam(pars, lr=1e-6, fused=True)     # one kernel for the 12 scalar parameters
    n4 = 1_000_000
    Px4, Py4 = random_pupil_device(n4, 0, dev)
    fields4 = lens4.get_field_coords()
    wls4 = list(lens4.wavelengths)
    S4 = lens4.surface_group.num_surfaces - 1
    last = [None]

    def c4():
        opt.zero_grad()
        last[0] = rms_merit_sharded(lens4, fields4, wls4, Px4, Py4, params=pars)
        opt.step()

    ms = timed(c4, max(3, args.steps // 2), 3)
    out["c4_cooke_adam_step_sharded_pairs"] = {
        "ms": ms, "ray_surfaces_per_s": 9 * n4 * S4 / (ms * 1e-3), "variables": len(pars),
        "pairs_per_rank": [len(parallel.shard_items(range(9), r, world)) for r in range(world)],
        "merit": float(last[0]),
        "api": "autograd.rms_merit_sharded: 9 (field, wavelength) pairs round-robin over the ranks, one FusedTraceFn "
               "node per rank (forward with blocked checkpoints + reduced-algebra adjoint), ONE all-reduce of the flat "
               "parameter gradient, Adam step on every rank"}
    return out


def reference_torch_eager():
    """The reference's own torch backend on this GPU (README.md:120 of the reference): C1 and C2,
    fp64 and fp32, 1 M rays per Optic.trace. Runs after this process has released its device memory."""
    from oracle import ref_bench
    out = {}
    if not ref_bench.available():
        return {"unavailable": "no reference on this box (oracle/_ref not built)"}
    for lens_name in ("c2", "c1"):
        for prec in ("float64", "float32"):
            key = f"{lens_name}_{prec}"
            try:
                r = ref_bench.run_subprocess(["torch", "--lens", lens_name, "--precision", prec, "--rays", 1_000_000,
                                              "--steps", 1, "--warmup", 1], timeout=600)
                out[key] = {"ray_surfaces_per_s": r["value"], "ms_per_step": r["ms_per_step"], "sample": r["sample"],
                            "rms_spot_mm": r["rms"]}
            except Exception as e:
                out[key] = {"error": repr(e)[:300]}
    return out


def run_extras(torch, lens, packed, npf, nf, S_eff, timed, args, fp32_peak, fp64_peak, hbm_peak=None):
    """Other precisions / modes on the same workload, one field at a time to bound memory."""
    from optiland_pr_b200 import flops
    from optiland_pr_b200.trace import fused_trace

    dev = torch.device("cuda", torch.cuda.current_device())
    out = {}
    rng = np.random.default_rng(7)
    n = npf * nf
    r_ = torch.from_numpy(rng.uniform(size=npf)).to(dev)
    th = torch.from_numpy(rng.uniform(0, 2 * np.pi, size=npf)).to(dev)
    Px, Py = torch.sqrt(r_) * torch.cos(th), torch.sqrt(r_) * torch.sin(th)
    Hx = torch.tensor([f[0] for f in FIELDS], device=dev).repeat_interleave(npf)
    Hy = torch.tensor([f[1] for f in FIELDS], device=dev).repeat_interleave(npf)
    rays = lens.generate_rays(Hx, Hy, Px.repeat(nf), Py.repeat(nf), WAVELENGTH)
    ins64 = rays.components()
    ins64[7] = None
    o64 = torch.empty((8, n), dtype=torch.float64, device=dev)
    fl_ray = flops.flops_per_ray(packed)
    for name, kw in (("fp64_replay_order_fma_fwd", {"fast": True, "replay": True}),):
        ms = timed(lambda: fused_trace(packed, ins64, out=o64, verify_nr=False, check_status=False, **kw), args.steps, 3)
        out[name] = {"ray_surfaces_per_s": n * S_eff / (ms * 1e-3), "ms": ms,
                     "frac_fp64_fma_peak": fl_ray * n / (ms * 1e-3) / 1e12 / fp64_peak}
    ins32 = [t.float() if t is not None else None for t in ins64]
    o32 = torch.empty((8, n), dtype=torch.float32, device=dev)
    ms = timed(lambda: fused_trace(packed, ins32, out=o32, verify_nr=False, check_status=False), args.steps, 3)
    out["fp32_fwd"] = {"ray_surfaces_per_s": n * S_eff / (ms * 1e-3), "ms": ms,
                       "frac_fp32_fma_peak": fl_ray * n / (ms * 1e-3) / 1e12 / fp32_peak,
                       "hbm_gbs": 60 * n / (ms * 1e-3) / 1e9}
    out["fp64_fma_peak_tflops"] = fp64_peak
    out["fp32_fma_peak_tflops"] = fp32_peak
    out.update(bench_other_configs(torch, timed, args, fp64_peak))
    try:
        from optiland_pr_b200.autograd import bench_backward
    except ImportError:
        bench_backward = None
    if bench_backward is not None:
        out.update(bench_backward(lens, packed, ins64, timed, args, fp64_peak, hbm_peak=hbm_peak))
    return out


def bench_other_configs(torch, timed, args, fp64_peak):
    """BASELINE configs C1 (Cooke triplet, 3 fields x 3 wavelengths x 1M rays, RMS spot per pair) and C3 (off-axis
    three-mirror polynomial freeform system: Newton-Raphson on every mirror, tilts, decenters -- the interpreter
    kernel) as throughput figures next to the headline; their VALUES are pinned by the parity tests."""
    from optiland_pr_b200 import samples, spot
    from optiland_pr_b200.trace import fused_trace
    out = {}
    dev = torch.device("cuda", torch.cuda.current_device())
    # ---- C1
    lens = samples.CookeTriplet()
    lens.surface_group.record = "none"
    n1 = 1_000_000
    rng = np.random.default_rng(0)
    r_, th = rng.uniform(size=n1), rng.uniform(0, 2 * np.pi, size=n1)
    Px = torch.from_numpy(np.sqrt(r_) * np.cos(th)).to(dev)
    Py = torch.from_numpy(np.sqrt(r_) * np.sin(th)).to(dev)
    S1 = lens.surface_group.num_surfaces - 1

    passes = [0]

    def c1():       # SpotDiagram: all 9 (field, wavelength) pairs in ONE fused launch; the radius pass re-reads kept states
        sd = spot.SpotDiagram(lens, distribution=(Px, Py))
        passes[0] = sd.passes
        return sd._rms.cpu()

    ms = timed(c1, max(2, args.steps // 2), 3)
    out["c1_cooke_9_pairs_x_1M"] = {"ms": ms, "ray_surfaces_per_s": 9 * n1 * S1 / (ms * 1e-3),
                                    "trace_passes": passes[0],
                                    "api": "spot.SpotDiagram(lens): 3 fields x 3 wavelengths in one fused launch + trace "
                                           "+ spot kernel (centroids), the RMS / geometric radii about the centroids from "
                                           "a streaming reduction of the kept final states, radii -> host"}
    # ---- C3 (BASELINE configs[2]: 64 M rays, seed 2): off-axis three-mirror system, 4 x 4 polynomial freeforms, rx tilts,
    # dy decenters -- the interpreter kernel. The pupil is drawn on the device (NumPy's stream for seed 2).
    from optiland_pr_b200 import flops
    from optiland_pr_b200.optic import random_pupil_device
    lens = samples.ThreeMirrorAnastigmat()
    wl = lens.primary_wavelength
    n3 = 1 << 26
    Px, Py = random_pupil_device(n3, 2, dev)
    f0 = lens.get_field_coords()[0]
    rays = lens.launch_rays(np.array([f0[0]]), np.array([f0[1]]), Px, Py, wl)
    ins = rays.components()
    ins[7] = None
    del Px, Py
    lens.surface_group.record = "none"
    packed = lens.surface_group.pack([float(wl)])
    o = torch.empty((8, n3), dtype=torch.float64, device=dev)
    res = fused_trace(packed, ins, out=o, fast=True)           # settles the Newton counts
    S3 = lens.surface_group.num_surfaces - 1
    fl3 = flops.flops_per_ray(packed, res.nr_iters)
    for name, fast in (("tolerance", True), ("strict", False)):
        ms = timed(lambda: fused_trace(packed, ins, out=o, fast=fast, verify_nr=False, check_status=False),
                   max(2, args.steps // 2), 2)
        out["c3_three_mirror_freeform_64M_" + name] = {
            "ms": ms, "rays": n3, "ray_surfaces_per_s": n3 * S3 / (ms * 1e-3), "nr_iters": res.nr_iters,
            "frac_fp64_fma_peak": fl3 * n3 / (ms * 1e-3) / 1e12 / fp64_peak, "flops_per_ray": fl3,
            "flop_model": flops.flop_model(packed, res.nr_iters),
            "kernel": "trace_fwd_kernel (interpreter: 4x4 polynomial freeforms, Newton-Raphson, tilts, decenters)"}
    del o, ins, rays
    torch.cuda.empty_cache()
    # ---- C4: one optimisation step through the public API (Optic.trace under autograd -> FusedTraceFn: forward with
    # tile-blocked checkpoints + reduced-algebra adjoint kernel), merit = sum over 3 wavelengths x 3 fields of the squared
    # RMS spot radius, variables = the six radii and six conic constants of the Cooke triplet (live torch Parameters), Adam update included
    lens4 = samples.CookeTriplet()
    lens4.surface_group.record = "none"
    sfs = lens4.surface_group.surfaces
    pars = []
    for k in range(1, 7):
        p_ = torch.nn.Parameter(torch.tensor(float(sfs[k].geometry.radius), dtype=torch.float64, device=dev))
        lens4.set_radius(p_, k)
        pars.append(p_)
    for k in range(1, 7):
        p_ = torch.nn.Parameter(torch.tensor(0.0, dtype=torch.float64, device=dev))
        lens4.set_conic(p_, k)
        pars.append(p_)
    opt = torch.optim.Adam(pars, lr=1e-6, fused=True)     # one kernel for the 12 scalar parameters
    n4 = 1_000_000
    r_, th = rng.uniform(size=n4), rng.uniform(0, 2 * np.pi, size=n4)
    Px4 = torch.from_numpy(np.sqrt(r_) * np.cos(th)).to(dev)
    Py4 = torch.from_numpy(np.sqrt(r_) * np.sin(th)).to(dev)
    fields4 = lens4.get_field_coords()
    wls4 = list(lens4.wavelengths)

    try:   # all 9 (wavelength, field) pairs as ONE autograd node (per-ray wavelengths)
        from optiland_pr_b200.autograd import _rms_merit_all

        def c4a():
            opt.zero_grad()
            loss = _rms_merit_all(lens4, fields4, wls4, Px4, Py4)
            loss.backward()
            opt.step()
            return loss

        ms = timed(c4a, max(2, args.steps // 2), 3)
        out["c4_cooke_merit_adam_step_one_node"] = {
            "ms": ms, "ray_surfaces_per_s": 9 * n4 * S1 / (ms * 1e-3), "variables": len(pars),
            "api": "generate_rays for all 9 pairs with per-ray wavelengths + SurfaceGroup.trace (one FusedTraceFn node) "
                   "+ spot.rms_spot_sq + loss.backward() + Adam step"}
    except Exception as e:   # an extra leg must never take the bench line down
        out["c4_cooke_merit_adam_step_one_node"] = {"error": repr(e)[:200]}
    # ---- C3, second half: refractive Zernike (fringe, 12 terms) + odd-asphere plate, tilts about three axes,
    # elliptical / rectangular / boolean apertures, TIR at the rim -- the interpreter kernel again, 64 M rays
    lens = samples.FreeformPlate()
    wl = lens.primary_wavelength
    nz = 1 << 26
    Px, Py = random_pupil_device(nz, 2, dev)
    f0 = lens.get_field_coords()[-1]     # the off-axis field: part of the pupil is past the critical angle
    rays = lens.launch_rays(np.array([f0[0]]), np.array([f0[1]]), Px, Py, wl)
    ins = rays.components()
    ins[7] = None
    del Px, Py
    lens.surface_group.record = "none"
    packed = lens.surface_group.pack([float(wl)])
    o = torch.empty((8, nz), dtype=torch.float64, device=dev)
    res = fused_trace(packed, ins, out=o, fast=True)
    Sz = lens.surface_group.num_surfaces - 1
    flz = flops.flops_per_ray(packed, res.nr_iters)
    ms = timed(lambda: fused_trace(packed, ins, out=o, fast=True, verify_nr=False, check_status=False),
               max(2, args.steps // 2), 2)
    out["c3_zernike_freeform_plate_64M_tolerance"] = {
        "ms": ms, "rays": nz, "ray_surfaces_per_s": nz * Sz / (ms * 1e-3), "nr_iters": res.nr_iters,
        "dead_rays": int(torch.isnan(o[0]).sum()), "clipped_rays": int((o[6] == 0).sum()),
        "frac_fp64_fma_peak": flz * nz / (ms * 1e-3) / 1e12 / fp64_peak, "flops_per_ray": flz,
        "flop_model": flops.flop_model(packed, res.nr_iters),
        "kernel": "trace_fwd_kernel (interpreter: Zernike fringe freeform through the merged tables, odd asphere, "
                  "aperture programs, tilted frames, TIR at the rim); flops_per_ray counts the reference's "
                  "term-by-term algorithm"}
    del o, ins, rays
    torch.cuda.empty_cache()
    return out


def bench_monte_carlo(torch, lens, S_eff, trials=4096, rays=262_144):
    """BASELINE config C5: 4096 perturbed copies of the lens x 262 144 rays each, merit = RMS spot at Hy = 1.
    One fused launch per round (olrt_trace_pupil_f64 with a prescription batch): no per-ray data in HBM."""
    from optiland_pr_b200.tolerancing import DistributionSampler, MonteCarloSpot, Perturbation
    s = lens.surface_group.surfaces
    curved = [k for k in range(1, len(s) - 1) if np.isfinite(float(s[k].geometry.radius))]
    perts = [Perturbation(lens, "radius", DistributionSampler("normal", seed=100 + k, loc=float(s[k].geometry.radius),
                                                              scale=1e-3 * abs(float(s[k].geometry.radius))), k)
             for k in curved]
    perts += [Perturbation(lens, "thickness", DistributionSampler("normal", seed=200 + k, loc=float(s[k].thickness),
                                                                  scale=0.02), k) for k in (1, 2, 3, 5, 8)]
    perts += [Perturbation(lens, "decenter", DistributionSampler("normal", seed=300 + k, loc=0.0, scale=0.1), k, axis=ax)
              for k, ax in ((1, "x"), (5, "y"), (7, "x"), (10, "y"))]
    rng = np.random.default_rng(3)
    r_, th = rng.uniform(size=rays), rng.uniform(0, 2 * np.pi, size=rays)
    Px = torch.from_numpy(np.sqrt(r_) * np.cos(th)).cuda()
    Py = torch.from_numpy(np.sqrt(r_) * np.sin(th)).cuda()
    mc = MonteCarloSpot(lens, perts)
    mc.run(64, [0.0], [1.0], WAVELENGTH, Px, Py)           # warm-up (workspace allocation)
    dts = []
    for _ in range(2):           # wall clock of a 0.2 s host + device job: the better of two runs
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        rms = mc.run(trials, [0.0], [1.0], WAVELENGTH, Px, Py)
        torch.cuda.synchronize()
        dts.append(time.perf_counter() - t0)
    dt = min(dts)
    t1 = time.perf_counter()
    MonteCarloSpot(lens, perts).draw(trials)
    t_draw = time.perf_counter() - t1
    rms = rms.cpu().numpy()
    return {"mc_tolerancing_c5": {
        "trials": trials, "rays_per_trial": rays, "perturbations": len(perts), "path": mc.last_path,
        "seconds": dt, "seconds_runs": dts, "seconds_host_draws": t_draw, "ray_surfaces_per_s": trials * rays * S_eff / dt,
        "rms_spot_mm_mean_std": [float(rms.mean()), float(rms.std())],
        "note": "wall clock of MonteCarloSpot.run: reference-order scalar draws on the host, nominal-lens Newton "
                "counts, chief rays, one fused launch per round, per-trial Newton-count check, merits read back"}}


if __name__ == "__main__":
    main()
