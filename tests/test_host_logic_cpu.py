"""Host-side logic that needs no GPU: which lenses the backward kernels cover (PackedSystem.differentiable mirrors
csrc/trace_bwd.cuh: bwd_supported), the kernel hint, and the scalar fast path of the packer."""
import numpy as np
import pytest
import torch

from optiland_pr_b200 import _lib as L
from optiland_pr_b200 import samples
from optiland_pr_b200.prescription import _f, pack_surface_group


@pytest.mark.parametrize("name,want,reduced", [
    ("CookeTriplet", True, True), ("DoubleGauss", True, True), ("EvenAsphereDoubleGauss", True, True),
    ("AsphericSinglet", True, True),
    ("ThreeMirrorAnastigmat", True, False), ("TiltedSinglet", True, False),   # general adjoint kernel only
    ("FreeformPlate", True, False),                                           # Zernike + odd asphere, tilted frames
    ("CoatedRelay", False, False), ("Anamorphic", False, False)])
def test_differentiable_set(name, want, reduced):
    lens = getattr(samples, name)()
    wl = [float(lens.wavelengths[lens.primary_index])]
    pk = pack_surface_group(lens.surface_group.surfaces, wl, record="all")
    assert pk.differentiable() is want
    assert pk.reduced_adjoint() is reduced
    if reduced:
        assert pk.fastkinds() and pk.make_sys().kernel_hint & L.HINT_FASTKINDS
        # every surface must be recorded: the records are the checkpoints
        assert not pack_surface_group(lens.surface_group.surfaces, wl, record="last").differentiable()


def test_scalar_fast_path_of_the_packer_keeps_the_value():
    for v in (1.5, -3, 0, float("inf"), np.float64(2.25), np.array([7.0]), torch.tensor(0.125, dtype=torch.float64)):
        got = _f(v)
        assert isinstance(got, float)
        assert got == float(np.asarray(v.detach() if hasattr(v, "detach") else v, dtype=np.float64).reshape(-1)[0])


def test_bx_limits_match_the_header():
    assert L.BX_MAXK == 3 and L.BWD_GENERIC == 1
