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


def test_prescription_tensor_chains_tilts_polynomial_and_zernike_leaves():
    """Host side of the extended differentiable set: a live tilt angle lands in the row as cos / sin of the angle, a
    live polynomial coefficient at its C[i][j] slot, a live Zernike coefficient twice (c and c N), each differentiable
    back to the leaf (autograd.prescription_tensor / _live_entries)."""
    from optiland_pr_b200.autograd import prescription_tensor
    # three-mirror system: rx of mirror 2, C[2][1] of mirror 1
    lens = samples.ThreeMirrorAnastigmat()
    sf = lens.surface_group.surfaces
    rx = torch.tensor(-0.174532, dtype=torch.float64, requires_grad=True)
    c21 = torch.tensor(2.97963986e-06, dtype=torch.float64, requires_grad=True)
    sf[2].geometry.cs.rx = rx
    co = [list(r) for r in sf[1].geometry.coefficients]
    co[2][1] = c21
    sf[1].geometry.coefficients = co
    pk = pack_surface_group(sf, [0.587], record="all")
    assert pk.differentiable() and not pk.reduced_adjoint()
    row = prescription_tensor(pk, sf, torch.float64, "cpu")
    assert np.array_equal(row.detach().numpy(), pk.presc[0])           # same values as the host pack
    lay2, lay1 = pk.layouts[1], pk.layouts[0]
    J = len(co[0])
    (row[lay2.frames + 7] * 2.0 + row[lay2.frames + 8] * 3.0 + row[lay1.coef + 2 * J + 1] * 5.0).backward()
    assert abs(float(rx.grad) - (-2.0 * np.sin(-0.174532) + 3.0 * np.cos(-0.174532))) < 1e-15
    assert float(c21.grad) == 5.0
    # Zernike plate (noll): coefficient 4 appears as c and c N
    plate = samples.FreeformPlate("noll")
    sp = plate.surface_group.surfaces
    z = sp[1].geometry.zernike
    cz = [torch.tensor(float(v), dtype=torch.float64, requires_grad=(j == 4)) if j == 4 else v
          for j, v in enumerate(z.coeffs)]
    z.coeffs = cz
    pk2 = pack_surface_group(sp, [0.55], record="all")
    assert pk2.differentiable() and pk2.coef_slots() == 24 and (pk2.make_sys().kernel_hint >> 8) == 24
    row2 = prescription_tensor(pk2, sp, torch.float64, "cpu")
    assert np.allclose(row2.detach().numpy(), pk2.presc[0], rtol=0, atol=0)
    n_, m_, off, norm = pk2.layouts[0].zernike_terms[4]
    (row2[off] * 7.0 + row2[off + 1] * 11.0).backward()
    assert abs(float(cz[4].grad) - (7.0 + 11.0 * norm)) < 1e-12 and norm != 1.0


@pytest.mark.parametrize("name, want", [("FreeformPlate", True), ("CookeTriplet", False), ("ThreeMirrorAnastigmat", False)])
def test_zernike_hint_is_announced_for_zernike_lenses_only(name, want):
    """olrt_sys.kernel_hint & OLRT_HINT_ZERNIKE (include/olrt.h): the ops table lives on the device, so the host says
    whether the launch should run the interpreter variant with the merged Zernike tables."""
    lens = getattr(samples, name)()
    packed = pack_surface_group(lens.surface_group.surfaces, [0.55])
    assert bool(packed.make_sys().kernel_hint & L.HINT_ZERNIKE) is want


@pytest.mark.parametrize("name", ["CookeTriplet", "EvenAsphereDoubleGauss", "FreeformPlate"])
def test_media_row_equals_a_full_pack_at_the_other_wavelength(name):
    """prescription.media_row: the row of the same surfaces at another wavelength is the packed row with the media
    entries (n_pre, n_post, k_pre, materials/base.py:73-119) re-evaluated -- bit for bit what a full pack gives."""
    from optiland_pr_b200.prescription import media_row
    lens = getattr(samples, name)()
    sf = lens.surface_group.surfaces
    p0 = pack_surface_group(sf, [0.55], record="none")
    for w in (0.4861, 0.6563):
        full = pack_surface_group(sf, [w], record="none")
        row = media_row(p0, sf, w)
        if not np.array_equal(full.ops, p0.ops):      # a medium absorbing at one wavelength only
            assert row is None
        else:
            assert row is not None and np.array_equal(row, full.presc[0])
    assert media_row(pack_surface_group(sf, [0.5, 0.6], record="none"), sf, 0.55) is None   # single-wavelength packs only
