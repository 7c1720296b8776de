"""optiland_pr_b200.dropin on the unmodified reference, the parts that need no GPU: with the patch installed and the
NumPy backend (or torch on the CPU) every call falls through to the stock loop and gives the stock bits; ParaxialRays
pass through; the packer and `prescription_tensor` read the reference's live torch objects (nn.Parameter-driven
radii / thicknesses / conics / asphere coefficients) and keep them in the autograd graph; the copy of the reference
under oracle/_ref is the reference."""
import hashlib
import os

import numpy as np
import pytest
import torch

import ref_cases as rc
from optiland_pr_b200 import dropin
from optiland_pr_b200.autograd import prescription_tensor
from optiland_pr_b200.prescription import pack_surface_group

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(autouse=True)
def _clean():
    yield
    dropin.uninstall()
    rc.restore()


def test_install_patches_only_surface_group_trace_and_uninstall_restores():
    from optiland.surfaces.surface_group import SurfaceGroup
    stock = SurfaceGroup.trace
    assert not dropin.installed()
    assert dropin.install() is stock
    assert dropin.installed() and SurfaceGroup.trace is not stock and SurfaceGroup.trace._olrt_fused
    assert dropin.install(fast_math=False) is stock       # idempotent: the stock method is remembered once
    dropin.uninstall()
    assert SurfaceGroup.trace is stock and not dropin.installed()


@pytest.mark.parametrize("backend", ["numpy", "torch-cpu"])
def test_off_gpu_every_call_is_the_stock_loop(backend):
    def run():
        if backend == "numpy":
            rc.use_numpy()
        else:
            rc.use_torch("cpu", "float64")
        lens = rc.make_lens("cooke")
        lens.paraxial.trace(1.0, 0.5, 0.55)                # ParaxialRays through SurfaceGroup.trace
        par = {"y": rc.to_np(lens.surface_group.y), "u": rc.to_np(lens.surface_group.u)}
        return rc.trace_records("cooke_stopped", 0.0, 1.0, 0.55, 6, "hexapolar"), rc.spot_diagram("cooke"), par
    want = run()
    dropin.install()
    got = run()
    st = dropin.stats()
    assert st.get("stock:not torch/cuda", 0) > 0 and not any(k.startswith("fused") for k in st)
    assert st.get("stock:not RealRays", 0) == 1           # paraxial_ray_tracer.py:56
    for k in want[0]:
        assert np.array_equal(got[0][k], want[0][k], equal_nan=True), k
    for k in want[1]:
        assert np.array_equal(got[1][k], want[1][k]), k
    assert want[2]["y"].shape[0] == 8 and np.array_equal(got[2]["y"], want[2]["y"])
    assert np.array_equal(got[2]["u"], want[2]["u"])


def test_packer_reads_live_reference_parameters_and_keeps_the_graph():
    """The reference under torch + grad mode: geometry.radius / .k / cs.z / coefficients are tensors computed from the
    optimiser's nn.Parameters (variable/*.py -> optic_updater.py). prescription_tensor must put exactly their values
    into the packed row and route d/d(row) back to the Parameters."""
    rc.use_torch("cpu", "float64", grad=True)
    lens = rc.make_lens("aspheric_singlet")
    p_r = torch.nn.Parameter(torch.tensor(20.5, dtype=torch.float64))
    p_k = torch.nn.Parameter(torch.tensor(-0.3, dtype=torch.float64))
    p_c = torch.nn.Parameter(torch.tensor(-2.0e-4, dtype=torch.float64))
    p_t = torch.nn.Parameter(torch.tensor(6.5, dtype=torch.float64))
    lens.set_radius(p_r * 1.0, 1)
    lens.set_conic(p_k * 1.0, 1)
    lens.set_asphere_coeff(p_c * 1.0, 1, 0)
    lens.set_thickness(p_t * 1.0, 1)
    surfaces = lens.surface_group.surfaces
    packed = pack_surface_group(surfaces, [0.55], record="all")
    assert packed.differentiable() and dropin._undifferentiated_variable(surfaces) is None
    row = prescription_tensor(packed, surfaces, torch.float64, torch.device("cpu"))
    assert np.array_equal(row.detach().numpy(), packed.presc[0])
    lay = packed.layouts[0]
    assert float(row[lay.radius]) == 20.5 and float(row[lay.conic]) == -0.3 and float(row[lay.coef]) == -2.0e-4
    wts = torch.arange(1, row.numel() + 1, dtype=torch.float64)
    (row * wts).sum().backward()
    assert float(p_r.grad) == float(wts[lay.radius]) and float(p_k.grad) == float(wts[lay.conic])
    assert float(p_c.grad) == float(wts[lay.coef])
    # the thickness after surface 1 moves cs.z of every later surface (optic_updater.py:67-85)
    later = [l.frames + 2 for l in packed.layouts[1:]]
    assert float(p_t.grad) == float(sum(wts[o] for o in later))


def test_variables_the_adjoint_does_not_differentiate_are_detected():
    rc.use_torch("cpu", "float64", grad=True)
    lens = rc.make_lens("cooke")
    surfaces = lens.surface_group.surfaces
    assert dropin._undifferentiated_variable(surfaces) is None      # grad-mode leaves are not variables
    p = torch.nn.Parameter(torch.tensor(0.01, dtype=torch.float64))
    surfaces[2].geometry.cs.rx = p * 1.0
    assert dropin._undifferentiated_variable(surfaces) is None      # tilts of a surface's own frame are differentiated
    surfaces[2].geometry.cs.rx = 0.0
    from optiland.coordinate_system import CoordinateSystem
    surfaces[2].geometry.cs.reference_cs = CoordinateSystem(rx=p * 1.0)
    assert "rx of a parent frame" in dropin._undifferentiated_variable(surfaces)
    surfaces[2].geometry.cs.reference_cs = None
    from optiland.materials import IdealMaterial
    mat = IdealMaterial(n=1.5)
    mat.index = (p * 0 + 1.5).reshape(1)
    surfaces[1].material_post = mat
    assert "refractive index" in dropin._undifferentiated_variable(surfaces)


def test_the_shipped_copy_is_the_reference():
    """oracle/_ref (git-ignored, travels to the GPU box) holds the reference's files verbatim."""
    manifest = os.path.join(ROOT, "oracle", "_ref", "MANIFEST.sha256")
    if not os.path.exists(manifest):
        pytest.skip("oracle/_ref not built (run __graft_entry__.build())")
    lines = open(manifest).read().split("\n")
    assert len(lines) > 200
    src = "/root/reference"
    checked = 0
    for ln in lines:
        if not ln:
            continue
        sha, rel = ln.split("  ", 1)
        with open(os.path.join(ROOT, "oracle", "_ref", rel), "rb") as fh:
            assert hashlib.sha256(fh.read()).hexdigest() == sha, rel
        if os.path.isdir(src):
            with open(os.path.join(src, rel), "rb") as fh:
                assert hashlib.sha256(fh.read()).hexdigest() == sha, rel
        checked += 1
    assert checked > 200
