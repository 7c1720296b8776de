"""The C-ABI library loads and exports every symbol include/olrt.h declares (no GPU needed)."""
import ctypes
import os
import re

import pytest

from optiland_pr_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "olrt.h")


def _header_text():
    with open(HEADER) as f:
        return f.read()


def _declared_functions():
    text = re.sub(r"/\*.*?\*/", "", _header_text(), flags=re.S)
    return sorted(set(re.findall(r"\b(olrt_[a-z0-9_]+)\s*\(", text)))


def _defines():
    out = {}
    for name, val in re.findall(r"#define\s+(OLRT_[A-Z0-9_]+)\s+\(?(-?\d+)u?\)?\s", _header_text()):
        out[name] = int(val)
    return out


def test_library_is_built():
    assert os.path.exists(_lib.LIB_PATH), "run `python -c 'import __graft_entry__ as g; g.build()'`"


def test_header_functions_match_binding_list():
    assert _declared_functions() == sorted(_lib.EXPORTS)


def test_every_declared_symbol_is_exported():
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in _declared_functions():
        assert hasattr(lib, name), name


def test_load_and_version():
    lib = _lib.load()
    assert lib.olrt_version() == _defines()["OLRT_ABI_VERSION"] == _lib.OLRT_ABI_VERSION
    assert b"sm_100a" in lib.olrt_build_info()


def test_constants_match_header():
    d = _defines()
    assert d["OLRT_MAX_SURFACES"] == _lib.MAX_SURFACES
    assert d["OLRT_MAX_WAVELENGTHS"] == _lib.MAX_WAVELENGTHS
    assert d["OLRT_MAX_NR_ITERS"] == _lib.MAX_NR_ITERS
    assert d["OLRT_OPS_STRIDE"] == _lib.OPS_STRIDE
    assert d["OLRT_SPOT_SUMS"] == _lib.SPOT_SUMS
    assert d["OLRT_FRAME_REALS"] == _lib.FRAME_REALS
    assert d["OLRT_GEOM_REALS"] == _lib.GEOM_REALS
    assert ctypes.sizeof(_lib.OlrtSys) == 8 * 4 + 8 + 8 * _lib.MAX_WAVELENGTHS


def test_argument_errors_do_not_need_a_gpu():
    """NULL / bad-size arguments are rejected before any CUDA call."""
    lib = _lib.load()
    sys_ = _lib.OlrtSys()
    arr = _lib.PtrArray8()
    rc = lib.olrt_trace_fwd_f64(ctypes.byref(sys_), None, None, 0, arr, None, arr, None, None, None, None, None, 0, None)
    assert rc == -1
    rc = lib.olrt_spot_sums_f64(None, None, None, 0, 0, 1, None, 0, None, None)
    assert rc == -1


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libolrt.so")
    with pytest.raises(_lib.OlrtLibraryError):
        _lib.load()
