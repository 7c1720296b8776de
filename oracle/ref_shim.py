"""TEST INFRASTRUCTURE ONLY -- import shim for the read-only reference.

Lets the unmodified reference at /root/reference (Optiland fork) be imported in
the build container, where matplotlib / vtk / seaborn are not installed. They
are imported at module scope on the hot path's import chain
(optiland/backend/numpy_backend.py:12, optiland/optic/optic.py:37-42) but never
called on it, so permissive stub modules are enough.

Used by oracle/make_golden*.py, by the CPU tests that pin the oracle against the live
reference, by the drop-in tests (tests/test_*dropin*.py: the fused trace under the
reference's own classes) and by bench.py's reference baselines. /root/reference does not
exist on the GPU box: there the verbatim copy under oracle/_ref/ (oracle/fetch_ref.py, made
by __graft_entry__.build(), git-ignored, shipped with the working tree) is imported
instead. The product package never imports this file.
"""
from __future__ import annotations

import importlib.machinery
import os
import sys
import types

_COPY = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


def _pick_reference_path() -> str:
    env = os.environ.get("OLRT_REFERENCE_PATH")
    if env:
        return env
    if os.path.isdir("/root/reference/optiland"):
        return "/root/reference"
    return _COPY


REFERENCE_PATH = _pick_reference_path()

_STUBS = [
    "matplotlib", "matplotlib.pyplot", "matplotlib.path", "matplotlib.axes",
    "matplotlib.figure", "matplotlib.patches", "matplotlib.colors", "matplotlib.cm",
    "matplotlib.ticker", "matplotlib.gridspec", "matplotlib.collections",
    "matplotlib.lines", "matplotlib.widgets", "matplotlib.backend_bases",
    "matplotlib.text", "matplotlib.transforms", "matplotlib.tri", "matplotlib.image",
    "matplotlib.animation", "matplotlib.backends", "matplotlib.backends.backend_agg",
    "mpl_toolkits", "mpl_toolkits.mplot3d", "mpl_toolkits.mplot3d.art3d",
    "mpl_toolkits.axes_grid1", "vtk", "seaborn",
]


class _Anything:
    """Callable / attribute sink used for every name looked up in a stub."""

    def __init__(self, *a, **k):
        pass

    def __call__(self, *a, **k):
        return _Anything()

    def __getattr__(self, name):
        if name.startswith("__") and name.endswith("__"):
            raise AttributeError(name)
        return _Anything()

    def __iter__(self):
        return iter(())

    def __mro_entries__(self, bases):
        return (object,)


class _StubModule(types.ModuleType):
    def __getattr__(self, name):
        if name.startswith("__") and name.endswith("__"):
            raise AttributeError(name)
        # classes so that `class Foo(matplotlib.X)` and isinstance() work
        cls = type(name, (_Anything,), {})
        setattr(self, name, cls)
        return cls


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_PATH, "optiland"))


def install() -> None:
    """Register the stubs and put the reference on sys.path (idempotent)."""
    if not reference_available():
        raise RuntimeError(f"reference not found at {REFERENCE_PATH}")
    for name in _STUBS:
        if name in sys.modules:
            continue
        try:
            if "." not in name:
                __import__(name)
                continue
        except Exception:
            pass
        mod = _StubModule(name)
        mod.__path__ = []  # behave as a package
        mod.__spec__ = importlib.machinery.ModuleSpec(name, None, is_package=True)
        sys.modules[name] = mod
        if "." in name:
            parent, child = name.rsplit(".", 1)
            if parent in sys.modules:
                setattr(sys.modules[parent], child, mod)
    if REFERENCE_PATH not in sys.path:
        sys.path.insert(0, REFERENCE_PATH)
