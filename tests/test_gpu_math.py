"""The kernels' own fp64 division / square root (csrc/math64.cuh) against IEEE results, bit for bit."""
import numpy as np
import pytest
import torch

from optiland_pr_b200 import _lib

pytestmark = pytest.mark.gpu


def _run(a, b):
    lib = _lib.load()
    a = torch.as_tensor(a, dtype=torch.float64, device="cuda").contiguous()
    b = torch.as_tensor(b, dtype=torch.float64, device="cuda").contiguous()
    q, r = torch.empty_like(a), torch.empty_like(a)
    rc = lib.olrt_selftest_math_f64(a.data_ptr(), b.data_ptr(), q.data_ptr(), r.data_ptr(), a.numel(),
                                    torch.cuda.current_stream().cuda_stream)
    _lib.check(rc, "selftest")
    torch.cuda.synchronize()
    return a.cpu().numpy(), b.cpu().numpy(), q.cpu().numpy(), r.cpu().numpy()


def _same(x, y, a=None, b=None):
    bad = (np.isnan(x) != np.isnan(y)) | (~np.isnan(y) & (x.view(np.int64) != y.view(np.int64)))
    if bad.any():
        k = np.nonzero(bad)[0][:8]
        print("mismatch at", k, "got", x[k], "want", y[k], "a", None if a is None else a[k],
              "b", None if b is None else b[k])
    return not bad.any()


@pytest.mark.parametrize("seed,lo,hi", [(0, -3, 3), (1, -30, 30), (2, -100, 100), (3, -0.01, 0.01)])
def test_random_operands_bit_exact(seed, lo, hi):
    rng = np.random.default_rng(seed)
    n = 1 << 22
    a = rng.uniform(1, 2, n) * 10.0 ** rng.uniform(lo, hi, n) * rng.choice([-1.0, 1.0], n)
    b = rng.uniform(1, 2, n) * 10.0 ** rng.uniform(lo, hi, n) * rng.choice([-1.0, 1.0], n)
    a, b, q, r = _run(a, b)
    with np.errstate(all="ignore"):
        assert _same(q, a / b)
        assert _same(r, np.sqrt(a))


def test_hard_mantissas_and_special_values():
    """Divisors with all-ones / near-power-of-two mantissas (the hard cases for a refined reciprocal),
    perfect squares, and the special values the trace can produce: 0, -0, inf, NaN, negative radicand."""
    ones = np.array([2.0 - 2.0**-52, 2.0 - 2.0**-51, 1.0 + 2.0**-52, 1.0 + 2.0**-51, 1.5, 1.0, 3.0, 7.0])
    rng = np.random.default_rng(9)
    num = np.concatenate([rng.uniform(1, 2, 4096), np.arange(1, 4097, dtype=np.float64)])
    a = np.repeat(num, ones.size)
    b = np.tile(ones, num.size)
    # magnitudes stay inside the documented domain of the fast path (2^-400 .. 2^400), math64.cuh
    sp = np.array([0.0, -0.0, 1.0, -1.0, np.inf, -np.inf, np.nan, 4.0, 1e-100, 1e100, 3.0e-7])
    a = np.concatenate([a, np.repeat(sp, sp.size), (np.arange(1, 2049, dtype=np.float64)) ** 2])
    b = np.concatenate([b, np.tile(sp, sp.size), np.ones(2048)])
    a, b, q, r = _run(a, b)
    with np.errstate(all="ignore"):
        ref_q, ref_r = a / b, np.sqrt(a)
    # signed zeros compare equal here: the fast path may return +0 where IEEE gives -0 (math64.cuh)
    nz = ref_q != 0
    assert _same(q[nz], ref_q[nz], a[nz], b[nz]) and np.all(q[~nz] == 0)
    nz = ref_r != 0
    assert _same(r[nz], ref_r[nz], a[nz]) and np.all(r[~nz] == 0)
