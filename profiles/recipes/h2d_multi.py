"""Concurrent pinned host->device bandwidth per rank (torchrun): is the multi-GPU e2e step bound by the host's PCIe complex?"""
import os, torch, torch.distributed as dist
local = int(os.environ.get("LOCAL_RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
torch.cuda.set_device(local)
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = 1 << 24
h = torch.empty(n, dtype=torch.float64).pin_memory(); h.fill_(1.0)
d = torch.empty(n, dtype=torch.float64, device="cuda")
for _ in range(3): d.copy_(h, non_blocking=True)
torch.cuda.synchronize()
if world > 1: dist.barrier()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): d.copy_(h, non_blocking=True)
e1.record(); torch.cuda.synchronize()
gbs = 20 * n * 8 / (e0.elapsed_time(e1) * 1e-3) / 1e9
t = torch.tensor([gbs], device="cuda")
if world > 1:
    parts = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(parts, t)
    if local == 0: print(f"world {world}: per-rank H2D GB/s", [round(float(p), 1) for p in parts], "sum", round(sum(float(p) for p in parts), 1))
    dist.destroy_process_group()
else:
    print(f"world 1: H2D GB/s {gbs:.1f}")
