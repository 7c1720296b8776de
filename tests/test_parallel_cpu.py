"""Host-side multi-GPU logic on CPU: world_size 2, gloo backend (the N > 1 path of bench.py / parallel.py)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from optiland_pr_b200 import parallel


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _sums(x, y, inten, shift):
    """olrt_spot_sums layout (include/olrt.h) for one group, all rays."""
    dx, dy = x - shift[0], y - shift[1]
    return np.array([[x.size, dx.sum(), dy.sum(), (dx * dx).sum(), (dy * dy).sum(), (dx * dx + dy * dy).max(),
                      inten.sum(), 0.0, shift[0], shift[1]]])


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(0)
    n = 10_001
    x, y, inten = rng.normal(size=n), rng.normal(size=n) + 3.0, rng.uniform(size=n)
    lo, hi = parallel.shard_range(n)
    shift = (x[0], y[0])                     # identical on every rank
    part = torch.tensor(_sums(x[lo:hi], y[lo:hi], inten[lo:hi], shift))
    merged = parallel.allreduce_spot_sums(part).numpy()
    whole = _sums(x, y, inten, shift)
    ok_sums = np.allclose(merged[:, :7], whole[:, :7], rtol=1e-13, atol=1e-12) and merged[0, 5] == whole[0, 5]
    iters = parallel.allreduce_nr_iters([1 + rank, 0, 2 - rank])
    g = torch.full((4,), float(rank + 1), dtype=torch.float64)
    parallel.allreduce_gradients(g)
    items = parallel.shard_items(range(9))
    # data-parallel parameter gradients: one all-reduce of the flat vector; a missing .grad counts as zero
    pa = torch.nn.Parameter(torch.zeros(3, dtype=torch.float64))
    pb = torch.nn.Parameter(torch.zeros((), dtype=torch.float64))
    pa.grad = torch.tensor([1.0, 2.0, 3.0], dtype=torch.float64) * (rank + 1)
    if rank == 1:
        pb.grad = torch.tensor(5.0, dtype=torch.float64)
    parallel.allreduce_param_grads([pa, pb])
    q.put((rank, (lo, hi), ok_sums, iters, g.tolist() + pa.grad.tolist() + [float(pb.grad)], items))
    dist.destroy_process_group()


def test_two_rank_reductions_on_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, rng0, ok0, it0, g0, items0), (r1, rng1, ok1, it1, g1, items1) = res
    assert rng0 == (0, 5001) and rng1 == (5001, 10001)          # contiguous, covering, first rank gets the extra
    assert ok0 and ok1
    assert it0 == it1 == [2, 0, 2]                                # element-wise max
    assert g0 == g1 == [3.0] * 4 + [3.0, 6.0, 9.0, 5.0]           # sums (plain vectors, then parameter gradients)
    assert items0 == [0, 2, 4, 6, 8] and items1 == [1, 3, 5, 7]


def test_shard_range_covers_everything():
    for n in (0, 1, 7, 8, 1000003):
        for w in (1, 2, 3, 8):
            edges = [parallel.shard_range(n, r, w) for r in range(w)]
            assert edges[0][0] == 0 and edges[-1][1] == n
            assert all(edges[r][1] == edges[r + 1][0] for r in range(w - 1))
            sizes = [hi - lo for lo, hi in edges]
            assert max(sizes) - min(sizes) <= 1
