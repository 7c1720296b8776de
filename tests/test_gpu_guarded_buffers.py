"""No kernel writes outside the buffers the host hands it: every device allocation of the host modules is fenced with
canary bytes (tests/guard_alloc.py) while each kernel family runs at RAGGED sizes (ray counts that are not multiples of
the warp, the CTA or the checkpoint tile). compute-sanitizer is closed on this GPU pool; this is the stand-in for its
memcheck on the write side."""
import numpy as np
import pytest
import torch

from optiland_pr_b200 import autograd as ag
from optiland_pr_b200 import backend as be
from optiland_pr_b200 import optic as optic_mod
from optiland_pr_b200 import rays as rays_mod
from optiland_pr_b200 import samples, spot, tolerancing, trace, wavefront
from optiland_pr_b200.optic import random_pupil_device
from optiland_pr_b200.prescription import pack_surface_group
from optiland_pr_b200.tolerancing import DistributionSampler, MonteCarloSpot, Perturbation

from guard_alloc import guarded
from helpers import load_golden, mirror_packed

pytestmark = pytest.mark.gpu
MODULES = [trace, ag, optic_mod, spot, tolerancing, wavefront, rays_mod]
RAGGED = (1, 31, 257, 4099, 70_001)


@pytest.fixture(autouse=True)
def _cuda():
    be.set_device("cuda:0")
    be.set_precision("float64")


def _golden_rays(name, n, dtype=torch.float64):
    g = load_golden(name)
    idx = np.arange(n) % g["rays_in"].shape[1]
    return g, [torch.tensor(g["rays_in"][q][idx], dtype=dtype, device="cuda") for q in range(8)]


@pytest.mark.parametrize("name", ["cooke_triplet", "even_asphere_double_gauss", "freeform_plate_fringe", "three_mirror",
                                  "anamorphic", "coated_relay"])
def test_array_mode_every_kernel_keeps_inside_its_buffers(name):
    with guarded(MODULES) as fence:
        for n in RAGGED[:4]:
            g, rays = _golden_rays(name, n)
            _, packed = mirror_packed(name, g["wavelengths"])        # record="all": surface records are written too
            for fast in (True, False):
                trace.fused_trace(packed, rays, fast=fast, want_dir0=True)
            trace.fused_trace(packed, [t.float() for t in rays], fast=True)
        torch.cuda.synchronize()
        assert len(fence.buffers) > 10
        assert fence.check() == []


def test_wild_rays_fixup_list_keeps_inside_its_buffers():
    lens = samples.EvenAsphereDoubleGauss()
    rng = np.random.default_rng(1)
    n = 20_011
    x, y = rng.uniform(-14, 14, n), rng.uniform(-14, 14, n)
    Lc, Mc = rng.uniform(-0.5, 0.5, n), rng.uniform(-0.5, 0.5, n)
    wild = np.stack([x, y, np.full(n, -8.0), Lc, Mc, np.sqrt(1 - Lc * Lc - Mc * Mc), np.ones(n), np.full(n, 0.5876)])
    packed = pack_surface_group(lens.surface_group.surfaces, [0.5876])
    with guarded(MODULES) as fence:
        r = trace.fused_trace(packed, [torch.tensor(wild[q], device="cuda") for q in range(8)], fast=True)
        torch.cuda.synchronize()
        assert r.fixups > 1000
        assert fence.check() == []


@pytest.mark.parametrize("npup", RAGGED[1:])
def test_pupil_and_group_mode_keep_inside_their_buffers(npup):
    lens = samples.EvenAsphereDoubleGauss()
    lens.surface_group.record = "none"
    px, py = random_pupil_device(max(npup, 8), 3, "cuda:0")
    px, py = px[:npup].contiguous(), py[:npup].contiguous()
    with guarded(MODULES) as fence:
        lens.trace([0.0, 0.0, 0.0], [0.0, 0.7, 1.0], 0.5876, distribution=(px, py), spot=True)
        s = lens.surface_group.surfaces
        mc = MonteCarloSpot(lens, [Perturbation(lens, "radius", DistributionSampler("normal", seed=1, loc=float(s[1].geometry.radius), scale=0.05), 1),
                                   Perturbation(lens, "decenter", DistributionSampler("normal", seed=2, loc=0.0, scale=0.1), 5, axis="y")])
        mc.run(5, [0.0, 0.0], [0.7, 1.0], 0.5876, px[:min(npup, 4099)], py[:min(npup, 4099)])
        torch.cuda.synchronize()
        assert fence.check() == []


def test_streamed_pupil_grid_keeps_inside_its_buffers():
    lens = samples.EvenAsphereDoubleGauss()
    lens.surface_group.record = "none"
    lens.pipeline_min_rays = 1 << 12
    lens.pipeline_chunks = 3
    px, py = random_pupil_device(3 * 65536 + 77, 4, "cuda:0")
    hx, hy = px.cpu().pin_memory(), py.cpu().pin_memory()
    with guarded(MODULES) as fence:
        lens.trace([0.0, 0.0], [0.0, 1.0], 0.5876, distribution=(hx, hy), spot=True)
        torch.cuda.synchronize()
        assert fence.check() == []


@pytest.mark.parametrize("n", RAGGED[1:4] + (9001,))
def test_forward_with_checkpoints_and_both_adjoints_keep_inside_their_buffers(n):
    lens = samples.EvenAsphereDoubleGauss()
    px, py = random_pupil_device(max(n, 8), 5, "cuda:0")
    rays = lens.generate_rays(torch.zeros(n, dtype=torch.float64, device="cuda"),
                              torch.full((n,), 0.7, dtype=torch.float64, device="cuda"), px[:n], py[:n], 0.5876)
    ins = rays.components()
    ins[7] = None
    with guarded(MODULES) as fence:
        for record, blocked, generic in (("none", True, False), ("all", False, False), ("all", False, True)):
            pk = pack_surface_group(lens.surface_group.surfaces, [0.5876], record=record)
            presc = torch.tensor(pk.presc[0], dtype=torch.float64, device="cuda", requires_grad=True)
            leaf = [t.detach().clone().requires_grad_(q < 6) if t is not None else None for q, t in enumerate(ins)]
            out, _, _ = ag.trace_differentiable(pk, presc, leaf, fast=True, blocked=blocked, generic_adjoint=generic)
            (out[0] ** 2 + out[1] ** 2).sum().backward()
            assert torch.isfinite(presc.grad).all()
        torch.cuda.synchronize()
        assert fence.check() == []


def test_opd_epilogue_and_spot_kernels_keep_inside_their_buffers():
    g = load_golden("wavefront_cooke")
    ck = samples.CookeTriplet()
    ck.set_paraxial(EPL=float(g["EPL"]), EPD=float(g["EPD"]), XPL=float(g["XPL"]))
    with guarded(MODULES) as fence:
        wavefront.Wavefront(ck, fields=[(0.0, 1.0)], wavelengths=[0.55], num_rays=7).get_data((0.0, 1.0), 0.55)
        sd = spot.SpotDiagram(ck, num_rings=5)
        sd.rms_spot_radius()
        torch.cuda.synchronize()
        assert fence.check() == []


def test_the_fence_sees_a_write_past_the_end():
    """The check itself: a deliberate one-element overrun is reported."""
    with guarded([trace]) as fence:
        t = trace.torch.empty(100, dtype=torch.float64, device="cuda")
        torch.as_strided(t, (101,), (1,))[100] = 1.0
        assert [lab for lab, _ in fence.check()] != []
