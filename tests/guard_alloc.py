"""Guarded device allocations for the out-of-bounds-write test (compute-sanitizer is closed on this GPU pool).

`guarded(modules)` swaps the name `torch` inside the given host modules for a proxy whose empty / zeros / empty_like /
zeros_like / full hand out CUDA tensors cut from the MIDDLE of a larger byte buffer filled with a canary pattern, PAD
bytes on either side. Every buffer the host side passes to a kernel (outputs, surface records, checkpoints, residual
words, fix-up lists, spot sums, gradient rows) is then fenced; `check()` asserts that no kernel wrote outside its
buffer. Reads out of bounds are not seen by this -- the parity tests at ragged sizes are the evidence for those."""
import contextlib
import math

import torch as _torch

PAD = 4096
CANARY = 0xA5


class _Fence:
    def __init__(self):
        self.buffers = []      # (raw uint8 buffer, nbytes, label)

    def alloc(self, shape, dtype, device, zero, label):
        if isinstance(shape, int):
            shape = (shape,)
        shape = tuple(int(v) for v in shape)
        item = _torch.empty((), dtype=dtype).element_size()
        nbytes = item * math.prod(shape)
        raw = _torch.full((nbytes + 2 * PAD,), CANARY, dtype=_torch.uint8, device=device)
        mid = raw[PAD:PAD + nbytes]
        if zero:
            mid.zero_()
        self.buffers.append((raw, nbytes, label))
        return mid.view(dtype).view(shape)

    def check(self):
        bad = []
        for raw, nbytes, label in self.buffers:
            if not bool((raw[:PAD] == CANARY).all()) or not bool((raw[PAD + nbytes:] == CANARY).all()):
                bad.append((label, nbytes))
        return bad


class _TorchProxy:
    def __init__(self, fence):
        self._fence = fence

    def __getattr__(self, name):
        return getattr(_torch, name)

    @staticmethod
    def _is_cuda(device):
        return device is not None and str(device).startswith("cuda")

    def _make(self, name, zero, *size, dtype=None, device=None, **kw):
        if not self._is_cuda(device) or kw.get("pin_memory") or kw.get("requires_grad"):
            return getattr(_torch, name)(*size, dtype=dtype, device=device, **kw)
        shape = size[0] if len(size) == 1 else size
        return self._fence.alloc(shape, dtype or _torch.get_default_dtype(), device, zero, f"{name}{tuple(size)}")

    def empty(self, *size, **kw):
        return self._make("empty", False, *size, **kw)

    def zeros(self, *size, **kw):
        return self._make("zeros", True, *size, **kw)

    def empty_like(self, t, **kw):
        if not t.is_cuda or kw:
            return _torch.empty_like(t, **kw)
        return self._fence.alloc(tuple(t.shape), t.dtype, t.device, False, "empty_like")

    def zeros_like(self, t, **kw):
        if not t.is_cuda or kw:
            return _torch.zeros_like(t, **kw)
        return self._fence.alloc(tuple(t.shape), t.dtype, t.device, True, "zeros_like")


@contextlib.contextmanager
def guarded(modules):
    fence = _Fence()
    proxy = _TorchProxy(fence)
    saved = [(m, m.torch) for m in modules]
    for m in modules:
        m.torch = proxy
    try:
        yield fence
    finally:
        for m, t in saved:
            m.torch = t
