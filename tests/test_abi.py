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
    # struct olrt_pupil (ABI v5): 5 pointers, 3 int64, 2 int32, 4 doubles, the ready pointer and its chunk length
    assert ctypes.sizeof(_lib.OlrtPupil) == 5 * 8 + 3 * 8 + 2 * 4 + 4 * 8 + 8 + 8
    assert d["OLRT_BWD_GENERIC"] == _lib.BWD_GENERIC
    assert d["OLRT_ST_TIMEOUT"] == _lib.ST_TIMEOUT


def test_argument_errors_do_not_need_a_gpu():
    """NULL / bad-size arguments are rejected before any CUDA call."""
    lib = _lib.load()
    sys_ = _lib.OlrtSys()
    arr = _lib.PtrArray8()
    rc = lib.olrt_trace_fwd_f64(ctypes.byref(sys_), None, None, 0, arr, None, arr, None, None, None, None, None, 0, None)
    assert rc == -1
    rc = lib.olrt_spot_sums_f64(None, None, None, 0, 0, 1, None, 0, None, None)
    assert rc == -1
    # streamed upload: NULL buffers and a chunk length that is not a multiple of 256 are refused before any copy
    assert lib.olrt_upload_pupil(None, None, None, None, 1024, 256, None, None, None) == -1
    buf = (ctypes.c_double * 4)()
    word = (ctypes.c_uint32 * 1)(1)
    a = ctypes.addressof(buf)
    assert lib.olrt_upload_pupil(a, a, a, a, 4, 100, ctypes.addressof(word), ctypes.addressof(word), None) == -1


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libolrt.so")
    with pytest.raises(_lib.OlrtLibraryError):
        _lib.load()
