import time, torch
x = torch.randn(1 << 26, device="cuda", dtype=torch.float64)
h = torch.empty(3, dtype=torch.float64).pin_memory()
def work():
    return (x[:50_000_000] * 1.0000001).sum().repeat(3)   # ~0.4 ms of GPU work
for name, fn in [
    ("copy_ blocking to pinned", lambda t: h.copy_(t)),
    ("copy_ non_blocking + stream sync", lambda t: (h.copy_(t, non_blocking=True), torch.cuda.current_stream().synchronize())),
    (".cpu()", lambda t: t.cpu()),
    (".tolist()", lambda t: t.tolist()),
    ("copy_ non_blocking + event sync", None),
]:
    for rep in range(3):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        t = work()
        if fn is None:
            h.copy_(t, non_blocking=True); e = torch.cuda.Event(); e.record(); e.synchronize()
        else:
            fn(t)
        t1 = time.perf_counter()
    print("%-36s %.3f ms" % (name, (t1 - t0) * 1e3))
