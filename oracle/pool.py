"""TEST INFRASTRUCTURE -- not part of the product. The NumPy oracle (oracle/trace_np.py) over a process pool, for
value-for-value comparisons at sizes one process cannot finish in seconds (tests/test_gpu_oracle_2m.py).

The reference's Newton exit test is a maximum over ALL rays of a batch (geometries/newton_raphson.py:148); chunks of a
batch therefore first report the trip counts their own rays need, the batch count is the maximum over chunks, and the
chunks that stopped earlier are traced again with the batch count -- exactly what one big call would have done."""
import multiprocessing as mp
import os

import numpy as np

from oracle import trace_np

_G = {}


def _init(ops, presc, wavelengths, max_iters):
    _G.update(ops=ops, presc=presc, wavelengths=wavelengths, max_iters=max_iters)


def _run(task):
    rays_in, fixed = task
    r = trace_np.Rays(*[a.copy() for a in rays_in])
    res = trace_np.trace(_G["ops"], _G["presc"], _G["wavelengths"], r, max_iters=_G["max_iters"], fixed_iters=fixed)
    return np.stack([r.x, r.y, r.z, r.L, r.M, r.N, r.i, r.opd]), list(res["nr_iters"])


def trace_batch(packed, rays_in, chunk=65_536, procs=None, presc_index=0):
    """rays_in: [8, n] float64 (x y z L M N intensity wavelength). Returns (out [8, n], batch nr_iters)."""
    n = rays_in.shape[1]
    procs = procs or min(16, os.cpu_count() or 1)
    cuts = list(range(0, n, chunk)) + [n]
    parts = [np.ascontiguousarray(rays_in[:, a:b]) for a, b in zip(cuts[:-1], cuts[1:])]
    max_iters = [lay.max_iter for lay in packed.layouts]
    ctx = mp.get_context("fork")     # the workers only ever touch NumPy
    with ctx.Pool(procs, initializer=_init,
                  initargs=(packed.ops, packed.presc[presc_index], list(packed.wavelengths), max_iters)) as pool:
        first = pool.map(_run, [(p, None) for p in parts])
        batch = [max(k[s] for _, k in first) for s in range(len(max_iters))]
        redo = [c for c, (_, k) in enumerate(first) if k != batch]
        again = pool.map(_run, [(parts[c], batch) for c in redo])
    outs = [o for o, _ in first]
    for c, (o, _) in zip(redo, again):
        outs[c] = o
    return np.concatenate(outs, axis=1), batch
