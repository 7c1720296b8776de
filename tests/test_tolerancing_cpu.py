"""CPU: host logic of the Monte-Carlo tolerancing driver (optiland_pr_b200/tolerancing.py): the perturbation draws
follow the reference's order and generators, and a batch row equals the prescription packed from the lens with
the same perturbations applied through the Optic updaters."""
import copy
import os

import numpy as np
import pytest

from oracle import ref_shim
from optiland_pr_b200 import samples
from optiland_pr_b200.prescription import pack_surface_group
from optiland_pr_b200.tolerancing import DistributionSampler, MonteCarloSpot, Perturbation, RangeSampler


def _perturbations(lens):
    s = lens.surface_group.surfaces
    return [
        Perturbation(lens, "radius", DistributionSampler("normal", seed=11, loc=float(s[1].geometry.radius),
                                                         scale=1e-3 * abs(float(s[1].geometry.radius))), 1),
        Perturbation(lens, "thickness", DistributionSampler("normal", seed=12, loc=float(s[2].thickness), scale=0.02), 2),
        Perturbation(lens, "decenter", DistributionSampler("normal", seed=13, loc=0.0, scale=0.1), 3, axis="x"),
        Perturbation(lens, "decenter", DistributionSampler("uniform", seed=14, low=-0.1, high=0.1), 5, axis="y"),
        Perturbation(lens, "conic", RangeSampler(-0.01, 0.01, 4), 5),
        Perturbation(lens, "asphere_coeff", DistributionSampler("normal", seed=15, loc=float(s[5].geometry.coefficients[1]),
                                                                scale=1e-9), 5, coeff_number=1),
    ]


def test_batch_rows_equal_packing_the_perturbed_lens():
    lens = samples.EvenAsphereDoubleGauss()
    mc = MonteCarloSpot(lens, _perturbations(lens))
    vals = mc.draw(7)
    packed = pack_surface_group(lens.surface_group.surfaces, [0.5876], record="none")
    batch = mc.prescriptions(packed, vals)
    assert batch.shape == (7, packed.presc.shape[1])
    for t in (0, 3, 6):
        pl = samples.EvenAsphereDoubleGauss()
        pl.set_radius(vals[t, 0], 1)
        pl.set_thickness(vals[t, 1], 2)
        pl.surface_group.surfaces[3].geometry.cs.x = vals[t, 2]
        pl.surface_group.surfaces[5].geometry.cs.y = vals[t, 3]
        pl.set_conic(vals[t, 4], 5)
        pl.set_asphere_coeff(vals[t, 5], 5, 1)
        want = pack_surface_group(pl.surface_group.surfaces, [0.5876], record="none")
        assert np.array_equal(want.ops, packed.ops)
        assert np.array_equal(batch[t], want.presc[0]), np.flatnonzero(batch[t] != want.presc[0])


def test_draw_order_is_trial_major_with_one_generator_per_perturbation():
    lens = samples.EvenAsphereDoubleGauss()
    mc = MonteCarloSpot(lens, _perturbations(lens))
    vals = mc.draw(5)
    g = np.random.default_rng(11)
    r1 = float(lens.surface_group.surfaces[1].geometry.radius)
    assert np.array_equal(vals[:, 0], [g.normal(loc=r1, scale=1e-3 * abs(r1)) for _ in range(5)])
    assert np.array_equal(vals[:, 4], [-0.01, np.linspace(-0.01, 0.01, 4)[1], np.linspace(-0.01, 0.01, 4)[2], 0.01, -0.01])


@pytest.mark.reference
@pytest.mark.skipif(not os.path.isdir(ref_shim.REFERENCE_PATH), reason="reference tree not present")
def test_samplers_match_the_reference_bit_for_bit():
    ref_shim.install()
    import optiland.backend as rbe
    rbe.set_backend("numpy")
    from optiland.tolerancing.perturbation import DistributionSampler as RefDist
    from optiland.tolerancing.perturbation import RangeSampler as RefRange
    for kw in (dict(distribution="normal", seed=3, loc=5.0, scale=0.25), dict(distribution="uniform", seed=4, low=-1.0, high=2.0)):
        a, b = DistributionSampler(**kw), RefDist(**kw)
        assert [a.sample() for _ in range(50)] == [float(b.sample()) for _ in range(50)]
    a, b = RangeSampler(0.0, 1.0, 7), RefRange(0.0, 1.0, 7)
    assert [a.sample() for _ in range(16)] == [float(b.sample()) for _ in range(16)]


def test_tilt_perturbation_rows_equal_packing_the_tilted_lens():
    """Tilt perturbations (cs.rx / ry := value) set the frame's cos / sin entries AND the frame flag of the ops table
    (the reference rotates only when the angle is truthy, coordinate_system.py:73-107)."""
    lens = samples.FreeformPlate()
    s = lens.surface_group.surfaces
    perts = [Perturbation(lens, "tilt", DistributionSampler("normal", seed=3, loc=float(s[3].geometry.cs.ry),
                                                            scale=0.003), 3, axis="y"),
             Perturbation(lens, "tilt", RangeSampler(0.0, 0.002, 3), 4, axis="x"),     # first value 0.0: flag stays off
             Perturbation(lens, "radius", DistributionSampler("normal", seed=4, loc=35.0, scale=0.02), 3)]
    mc = MonteCarloSpot(lens, perts)
    vals = mc.draw(3)
    packed = pack_surface_group(s, [0.55], record="none")
    with pytest.raises(NotImplementedError):
        mc.prescriptions(packed, vals)
    ops_rows = []
    batch = mc.prescriptions(packed, vals, ops_out=ops_rows)
    assert len(ops_rows) == 3
    for t in range(3):
        pl = samples.FreeformPlate()
        ps = pl.surface_group.surfaces
        ps[3].geometry.cs.ry = float(vals[t, 0])
        ps[4].geometry.cs.rx = float(vals[t, 1])
        ps[3].geometry.radius = float(vals[t, 2])
        want = pack_surface_group(ps, [0.55], record="none")
        assert np.array_equal(batch[t], want.presc[0]), t
        assert np.array_equal(ops_rows[t], want.ops), t
    assert vals[0, 1] == 0.0 and not np.array_equal(ops_rows[0], ops_rows[1])


def test_vectorised_draws_equal_scalar_draws():
    """MonteCarloSpot draws every perturbation's values in one call per sampler: same values, same generator state."""
    for kind, kw in (("normal", dict(loc=3.0, scale=0.1)), ("uniform", dict(low=-1.0, high=2.0))):
        a, b = DistributionSampler(kind, seed=5, **kw), DistributionSampler(kind, seed=5, **kw)
        x = np.array([a.sample() for _ in range(500)])
        y = np.concatenate([b.sample_n(123), b.sample_n(377)])
        assert np.array_equal(x, y) and a.sample() == b.sample()
    a, b = RangeSampler(-1, 1, 7), RangeSampler(-1, 1, 7)
    x = np.array([a.sample() for _ in range(20)])
    y = np.concatenate([b.sample_n(5), b.sample_n(15)])
    assert np.array_equal(x, y) and a.sample() == b.sample()
