"""GPU, two ranks: BASELINE C4 sharded over (field, wavelength) pairs (autograd.rms_merit_sharded) gives the same
merit and the same parameter gradients on every rank as one rank tracing all nine pairs, to 1e-12; and the
Monte-Carlo trials of C5 sharded over ranks (MonteCarloSpot.run) give the single-rank merits.
Both ranks use cuda:0 (the driver's GPU test box has one GPU) and exchange their few hundred bytes over gloo; on a
multi-GPU box bench.py runs the same code over NCCL."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _problem(dev):
    from optiland_pr_b200 import backend as be
    from optiland_pr_b200 import samples
    be.set_device(str(dev))
    be.set_precision("float64")
    lens = samples.CookeTriplet()
    lens.surface_group.record = "none"
    sfs = lens.surface_group.surfaces
    pars = []
    for k in range(1, 7):
        p_ = torch.nn.Parameter(torch.tensor(float(sfs[k].geometry.radius), dtype=torch.float64, device=dev))
        lens.set_radius(p_, k)
        pars.append(p_)
    for k in (1, 3, 5):
        p_ = torch.nn.Parameter(torch.tensor(float(sfs[k].thickness), dtype=torch.float64, device=dev))
        lens.set_thickness(p_, k)
        pars.append(p_)
    rng = np.random.default_rng(3)
    r_, th = rng.uniform(size=20_000), rng.uniform(0, 2 * np.pi, size=20_000)
    Px = torch.from_numpy(np.sqrt(r_) * np.cos(th)).to(dev)
    Py = torch.from_numpy(np.sqrt(r_) * np.sin(th)).to(dev)
    return lens, pars, Px, Py


def _mc(dev):
    from optiland_pr_b200 import samples
    from optiland_pr_b200.tolerancing import DistributionSampler, MonteCarloSpot, Perturbation
    lens = samples.EvenAsphereDoubleGauss()
    s = lens.surface_group.surfaces
    perts = [Perturbation(lens, "radius", DistributionSampler("normal", seed=10 + k, loc=float(s[k].geometry.radius),
                                                              scale=1e-3 * abs(float(s[k].geometry.radius))), k)
             for k in (1, 5, 9)]
    perts += [Perturbation(lens, "decenter", DistributionSampler("normal", seed=30, loc=0.0, scale=0.1), 7, axis="x")]
    rng = np.random.default_rng(4)
    r_, th = rng.uniform(size=3000), rng.uniform(0, 2 * np.pi, size=3000)
    return MonteCarloSpot(lens, perts), np.sqrt(r_) * np.cos(th), np.sqrt(r_) * np.sin(th)


def _worker(rank, world, port, q):
    import torch.distributed as dist
    from optiland_pr_b200.autograd import rms_merit_sharded
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    lens, pars, Px, Py = _problem(dev)
    loss = rms_merit_sharded(lens, lens.get_field_coords(), list(lens.wavelengths), Px, Py, params=pars)
    mc, px, py = _mc(dev)
    rms = mc.run(7, [0.0, 0.0], [0.7, 1.0], 0.5876, px, py).cpu().numpy()
    q.put((rank, float(loss), [float(p.grad) for p in pars], rms))
    dist.destroy_process_group()


def test_two_ranks_equal_one_rank():
    from optiland_pr_b200.autograd import rms_merit_sharded
    dev = torch.device("cuda", 0)
    lens, pars, Px, Py = _problem(dev)
    want_loss = float(rms_merit_sharded(lens, lens.get_field_coords(), list(lens.wavelengths), Px, Py, params=pars))
    want_g = np.array([float(p.grad) for p in pars])
    mc, px, py = _mc(dev)
    want_rms = mc.run(7, [0.0, 0.0], [0.7, 1.0], 0.5876, px, py).cpu().numpy()
    assert np.isfinite(want_g).all() and (want_g != 0).all()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted((q.get(timeout=300) for _ in procs), key=lambda t: t[0])
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    for rank, loss, g, rms in res:
        assert abs(loss - want_loss) <= 1e-12 * abs(want_loss), (rank, loss, want_loss)
        assert np.max(np.abs(np.array(g) - want_g) / np.abs(want_g)) <= 1e-12, rank
        assert rms.shape == (7, 2) and np.allclose(rms, want_rms, rtol=1e-13, atol=0), rank
