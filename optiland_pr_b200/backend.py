"""Backend configuration mirroring the part of `optiland.backend` this path reads.

    be.set_backend / get_backend        optiland/backend/__init__.py:95-120
    set_device / set_precision / grad_mode   optiland/backend/torch_backend.py:45-127

Only the torch backend on CUDA exists here: the package is a B200 accelerator, not a second
NumPy implementation. `set_backend("numpy")` raises instead of falling back.
"""
from __future__ import annotations

import contextlib

import torch

pi = 3.141592653589793
inf = float("inf")
nan = float("nan")


class GradMode:
    """torch_backend.py:33-64"""

    def __init__(self):
        self.requires_grad = False

    def enable(self):
        self.requires_grad = True

    def disable(self):
        self.requires_grad = False

    @contextlib.contextmanager
    def temporary_enable(self):
        old = self.requires_grad
        self.requires_grad = True
        try:
            yield
        finally:
            self.requires_grad = old


class _Config:
    def __init__(self):
        self.backend = "torch"
        self.device = "cuda"
        self.precision = torch.float64
        self.grad_mode = GradMode()


_config = _Config()
grad_mode = _config.grad_mode


def set_backend(name: str) -> None:
    if name != "torch":
        raise ValueError(
            f"optiland_pr_b200 only implements the torch/CUDA backend (got {name!r}); "
            "use the reference package for the NumPy backend")
    _config.backend = name


def get_backend() -> str:
    return _config.backend


def list_available_backends():
    return ["torch"]


def set_device(device: str) -> None:
    if not str(device).startswith("cuda"):
        raise ValueError("optiland_pr_b200 has no CPU path; device must be 'cuda' or 'cuda:N'")
    _config.device = device


def get_device() -> str:
    return _config.device


def set_precision(precision: str) -> None:
    if precision == "float32":
        _config.precision = torch.float32
    elif precision == "float64":
        _config.precision = torch.float64
    else:
        raise ValueError("Precision must be 'float32' or 'float64'.")


def get_precision() -> torch.dtype:
    return _config.precision


def array(x):
    """torch_backend.py:131-148"""
    if isinstance(x, torch.Tensor):
        return x
    return torch.tensor(x, device=get_device(), dtype=get_precision(),
                        requires_grad=grad_mode.requires_grad)


def as_array_1d(x):
    """Force a 1-D tensor (scalars become length-1), like optiland.backend.as_array_1d."""
    if isinstance(x, torch.Tensor):
        t = x.to(device=get_device(), dtype=get_precision())
    else:
        t = torch.as_tensor(x, dtype=get_precision(), device=get_device())
    return t.reshape(-1)


def to_numpy(x):
    if isinstance(x, torch.Tensor):
        return x.detach().cpu().numpy()
    import numpy as np

    return np.asarray(x)


def empty(n):
    return torch.empty(n, device=get_device(), dtype=get_precision())


def size(x) -> int:
    return int(x.numel()) if isinstance(x, torch.Tensor) else int(len(x))


def stack(xs, dim=0):
    return torch.stack(list(xs), dim=dim)
