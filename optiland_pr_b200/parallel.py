"""Multi-GPU plumbing: one process per GPU (torchrun), `torch.distributed` over NCCL (gloo in the CPU tests).

The path shards trivially: rays are independent, so ranks take contiguous ray ranges of
a big batch, or whole (field, wavelength) pairs / Monte-Carlo trials, with the prescription replicated.
No ray data ever crosses NVLink. The only exchanges are scalars:

  * spot sums   [G][10] doubles per merit evaluation   -> SUM (moments) / MAX (geometric radius)
  * Newton trip counts K_s, one int per NR surface     -> MAX (the reference's exit test is a maximum
    over ALL rays of the batch, newton_raphson.py:148, so the shards must agree on it)
  * parameter gradients [stride] doubles per backward  -> SUM

All are <= ~1 KB: latency-bound collectives; they are issued on the current stream right after the
kernel that produced them.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def world():
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def bind_to_gpu_numa(local_rank: int) -> list | None:
    """Pin this process to the CPUs closest to GPU `local_rank` (NVML's ideal affinity: the GPU's NUMA node) BEFORE it
    allocates pinned host buffers, so that the host -> device stream of a rank does not cross the socket
    interconnect. One process per GPU makes this the rank's whole affinity. Returns the CPU list, or None when NVML
    or the affinity call is unavailable (nothing is changed then)."""
    import os
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local_rank)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = [64 * w + b for w, word in enumerate(words) for b in range(64) if (word >> b) & 1]
        cpus = [c for c in cpus if c in os.sched_getaffinity(0)]
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return cpus
    except Exception:
        return None


def shard_range(n: int, rank: int | None = None, world_size: int | None = None):
    """Contiguous [lo, hi) of n items for this rank; the first n % world ranks get one extra."""
    r, w = world()
    rank = r if rank is None else rank
    world_size = w if world_size is None else world_size
    base, extra = divmod(n, world_size)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_items(items, rank: int | None = None, world_size: int | None = None):
    """Round-robin assignment of whole work items ((field, wavelength) pairs, Monte-Carlo trials)."""
    r, w = world()
    rank = r if rank is None else rank
    world_size = w if world_size is None else world_size
    return list(items)[rank::world_size]


def allreduce_spot_sums(sums: torch.Tensor) -> torch.Tensor:
    """Merge per-rank spot sums (layout include/olrt.h: olrt_spot_sums_*). Every rank must have used the
    same shift (columns 8, 9) -- pass an explicit `shift` to spot_sums when sharding rays of one group.
    Columns 0-4 and 6 add; column 5 (max r^2, a non-negative double, NaN marks a NaN ray) takes the max.
    ONE collective: the [G, 10] rows of every rank are gathered (< 1 KB each) and merged locally."""
    _, w = world()
    if w == 1:
        return sums
    sums = sums.contiguous()
    parts = [torch.empty_like(sums) for _ in range(w)]
    dist.all_gather(parts, sums)
    buf = torch.stack(parts)
    out = buf.sum(dim=0)
    out[:, 5] = buf[:, :, 5].amax(dim=0)      # amax propagates NaN
    out[:, 7:] = sums[:, 7:]
    return out


def allreduce_nr_iters(iters, device=None):
    """Agree on the Newton trip counts: element-wise MAX over ranks."""
    _, w = world()
    if w == 1:
        return list(iters)
    t = torch.tensor(list(iters), dtype=torch.int32, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [int(v) for v in t.cpu().tolist()]


def allreduce_gradients(*grads: torch.Tensor):
    """SUM the (tiny) parameter-gradient vectors over ranks, in place."""
    _, w = world()
    if w > 1:
        for g in grads:
            if g is not None:
                dist.all_reduce(g, op=dist.ReduceOp.SUM)
    return grads


def allreduce_param_grads(params):
    """Data-parallel gradient exchange for an optimiser step: ONE all-reduce (SUM) of the flat vector of every
    parameter's gradient (a missing .grad counts as zero), written back in place. A few hundred bytes."""
    _, w = world()
    params = [p for p in params]
    if w == 1 or not params:
        return
    dev, dt = params[0].device, params[0].dtype
    flat = torch.cat([(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1).to(dt) for p in params])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM)
    off = 0
    for p in params:
        k = p.numel()
        g = flat[off:off + k].reshape(p.shape).to(p.dtype)
        if p.grad is None:
            p.grad = g.clone()
        else:
            p.grad.copy_(g)
        off += k


def sharded_trace(surface_group, rays, wavelengths=None):
    """`SurfaceGroup.trace` on this rank's shard with BATCH-global Newton trip counts: trace the shard
    (its own exit rule), all-reduce K_s with MAX -- the reference's test is a maximum over all rays of the
    batch (newton_raphson.py:148) -- and re-trace with the agreed counts where this shard alone needed fewer."""
    _, w = world()
    names = ("x", "y", "z", "L", "M", "N", "i", "opd")
    start = [getattr(rays, q) for q in names]  # trace() replaces these attributes, it does not write in place
    out = surface_group.trace(rays, wavelengths=wavelengths)
    if w == 1:
        return out
    local = list(surface_group.last_trace.nr_iters)
    agreed = allreduce_nr_iters(local, device=rays.x.device)
    if agreed != local:
        surfaces = [s for s in surface_group.surfaces if type(s).__name__ != "ObjectSurface"]
        surface_group.forced_nr_iters = {surface_group.surfaces.index(s): k for s, k in zip(surfaces, agreed)}
        for q, v in zip(names, start):
            setattr(rays, q, v)
        try:
            out = surface_group.trace(rays, wavelengths=wavelengths)
        finally:
            surface_group.forced_nr_iters = None
    return out
