"""ctypes binding of libolrt.so (include/olrt.h).

The library is built in-tree (`optiland_pr_b200/libolrt.so`, see csrc/Makefile and
`__graft_entry__.build`). There is NO fallback: if the shared object is missing or a
symbol is absent, loading raises -- the product path never silently runs on the CPU.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("OLRT_LIB") or os.path.join(_HERE, "libolrt.so")  # OLRT_LIB: A/B builds only

OLRT_ABI_VERSION = 6
MAX_SURFACES = 64
MAX_WAVELENGTHS = 8
MAX_NR_ITERS = 128
OPS_STRIDE = 16
RESID_STRIDE = MAX_NR_ITERS + 1
SPOT_SUMS = 10
GROUP_RESID = 8
NOALLOC_MAX_RAYS = 1 << 20   # olrt_trace_fwd_noalloc_f64: rays per call (its re-trace list is a static device array)
FRAME_REALS = 9
GEOM_REALS = 4

# ops row fields
(OP_GEOM, OP_INTERACT, OP_FLAGS, OP_DOFF, OP_NCOEF, OP_AUX, OP_NR_ITERS, OP_APER_OFF,
 OP_APER_LEN, OP_RECORD, OP_NFRAMES, OP_FRAMEFLAGS, OP_APER_DOFF) = range(13)

GEOM_PLANE, GEOM_STANDARD, GEOM_EVEN_ASPHERE, GEOM_POLYNOMIAL, GEOM_ZERNIKE, GEOM_ODD_ASPHERE = range(6)
GEOM_BICONIC, GEOM_TOROIDAL, GEOM_CHEBYSHEV = 6, 7, 8
F_RADIUS_INF, F_ABSORB, F_COATING, F_THIN_LENS, F_RENORM = 1, 2, 4, 8, 16
AP_RADIAL, AP_OFFSET_RADIAL, AP_ELLIPTICAL, AP_RECTANGULAR = 1, 2, 3, 4
AP_OR, AP_AND, AP_DIFF = 10, 11, 12
ST_ZERNIKE_RANGE, ST_WAVELENGTH, ST_BAD_HINT, ST_CHEBYSHEV_RANGE, ST_TIMEOUT = 1, 2, 4, 8, 16
HINT_FASTKINDS = 1
HINT_ORDINARY = 2          # centred, refracting, aperture-free surfaces only (include/olrt.h)
HINT_ZERNIKE = 4           # a Zernike surface: the tolerance-mode interpreter variant with merged term tables
FWD_CKPT_BLOCKED = 4     # olrt_trace_fwd_* flag: `record` is the tile-blocked checkpoint buffer (include/olrt.h)
BWD_CKPT_BLOCKED = 2     # olrt_trace_bwd_* flag: `record` is that buffer
BWD_NO_FRAME_XY = 4      # olrt_trace_bwd_* flag: d loss / d (frame x, frame y) is not wanted (no decenter variable)
CKPT_TILE = 256
CKPT_VALS = 6            # x y z L M copysign(intensity, N) per ray and surface (include/olrt.h)
BWD_GENERIC = 1          # olrt_trace_bwd_* flag: force the general adjoint kernel (csrc/trace_bwd.cuh)
BX_MAXK = 3              # Newton updates per surface the reduced-algebra adjoint kernel (csrc/trace_bx.cuh) unrolls
BX_MAXC = 8              # even-asphere coefficients per surface in the reduced adjoint (OLRT_BWD_MAXC)
BWG_MAXC = 32            # coefficient gradient slots per surface in the general adjoint (OLRT_BWG_MAXC)
BWG_MAXM = 12            # largest |m| of a Zernike term there (OLRT_BWG_MAXM)
FWD_FAST = 1
FWD_REPLAY = 2
SPOT_VALID_ONLY = 1

ERRORS = {0: "ok", -1: "OLRT_E_ARG", -2: "OLRT_E_LIMIT", -3: "OLRT_E_LAUNCH", -4: "OLRT_E_UNSUPPORTED"}


class OlrtSys(C.Structure):
    """struct olrt_sys (include/olrt.h)."""

    _fields_ = [
        ("n_surfaces", C.c_int32),
        ("n_wavelengths", C.c_int32),
        ("n_presc", C.c_int32),
        ("presc_stride", C.c_int32),
        ("ops_len", C.c_int32),
        ("n_record", C.c_int32),
        ("nr_resid_words", C.c_int32),
        ("kernel_hint", C.c_int32),
        ("rays_per_presc", C.c_int64),
        ("wavelengths", C.c_double * MAX_WAVELENGTHS),
    ]


class OlrtPupil(C.Structure):
    """struct olrt_pupil (include/olrt.h): in-kernel ray launch for olrt_trace_pupil_*."""

    _fields_ = [
        ("Px", C.c_void_p), ("Py", C.c_void_p), ("field_xy", C.c_void_p),
        ("field_vig", C.c_void_p), ("field_z", C.c_void_p),
        ("n_pupil", C.c_int64), ("p_lo", C.c_int64), ("p_len", C.c_int64),
        ("n_fields", C.c_int32), ("reserved0", C.c_int32),
        ("EPD", C.c_double), ("EPL", C.c_double), ("z_start", C.c_double), ("wavelength", C.c_double),
        ("ready", C.c_void_p), ("ready_chunk", C.c_int64),
    ]


class OlrtRefSphere(C.Structure):
    """struct olrt_refsphere (include/olrt.h)."""

    _fields_ = [(k, C.c_double) for k in ("xc", "yc", "zc", "R", "n_image", "opd_ref", "wavelength", "tilt_ux",
                                          "tilt_uy", "epd")]


OPD_RAW = 1
PtrArray8 = C.c_void_p * 8

# every symbol include/olrt.h declares (tests/test_abi.py checks the header against this list)
EXPORTS = [
    "olrt_version", "olrt_build_info",
    "olrt_trace_fwd_f64", "olrt_trace_fwd_f32", "olrt_trace_fwd_noalloc_f64",
    "olrt_trace_bwd_f64", "olrt_trace_bwd_f32",
    "olrt_spot_sums_f64", "olrt_spot_sums_f32",
    "olrt_fma_peak_f64", "olrt_fma_peak_f32",
    "olrt_selftest_math_f64",
    "olrt_generate_rays_f64", "olrt_generate_rays_f32",
    "olrt_trace_pupil_f64", "olrt_trace_pupil_f32",
    "olrt_opd_epilogue_f64", "olrt_opd_epilogue_f32",
    "olrt_upload_pupil", "olrt_pupil_random_f64", "olrt_pupil_hexapolar_f64", "olrt_pupil_uniform_f64",
]

_lib = None


class OlrtLibraryError(RuntimeError):
    pass


def load():
    """dlopen libolrt.so and declare the prototypes. Raises if it is not there."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise OlrtLibraryError(
            f"{LIB_PATH} not found: build it with `make -C optiland_pr_b200/csrc` "
            "(or `python -c 'import __graft_entry__ as g; g.build()'`). There is no CPU fallback."
        )
    lib = C.CDLL(LIB_PATH)
    for name in EXPORTS:
        if not hasattr(lib, name):
            raise OlrtLibraryError(f"libolrt.so does not export {name}")
    lib.olrt_version.restype = C.c_int
    lib.olrt_upload_pupil.restype = C.c_int
    lib.olrt_upload_pupil.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64,
                                      C.c_void_p, C.c_void_p, C.c_void_p]
    lib.olrt_build_info.restype = C.c_char_p
    v = C.c_void_p
    fwd_args = [C.POINTER(OlrtSys), v, v, C.c_int64, PtrArray8, v, PtrArray8, v, v, v, v, v, C.c_int, v]
    bwd_args = [C.POINTER(OlrtSys), v, v, C.c_int64, PtrArray8, v, PtrArray8, v, PtrArray8, v, v, C.c_int, v]
    for suf in ("f64", "f32"):
        f = getattr(lib, f"olrt_trace_fwd_{suf}")
        f.restype, f.argtypes = C.c_int, fwd_args
        f = getattr(lib, f"olrt_trace_bwd_{suf}")
        f.restype, f.argtypes = C.c_int, bwd_args
        f = getattr(lib, f"olrt_spot_sums_{suf}")
        f.restype = C.c_int
        f.argtypes = [v, v, v, C.c_int64, C.c_int64, C.c_int32, v, C.c_int, v, v]
        f = getattr(lib, f"olrt_fma_peak_{suf}")
        f.restype, f.argtypes = C.c_int, [v, C.c_int64, C.c_int, C.c_int, v]
    lib.olrt_trace_fwd_noalloc_f64.restype = C.c_int   # olrt_trace_fwd_f64 without the `fixup` argument
    lib.olrt_trace_fwd_noalloc_f64.argtypes = fwd_args[:11] + fwd_args[12:]
    for suf in ("f64", "f32"):
        f = getattr(lib, f"olrt_generate_rays_{suf}")
        f.restype = C.c_int
        f.argtypes = [C.POINTER(OlrtPupil), PtrArray8, v]
    for suf in ("f64", "f32"):
        f = getattr(lib, f"olrt_trace_pupil_{suf}")
        f.restype = C.c_int
        f.argtypes = [C.POINTER(OlrtSys), v, v, C.POINTER(OlrtPupil), C.POINTER(PtrArray8), v, v, v, v, v, v,
                      C.c_int, v, v, C.c_int, v]
    for suf in ("f64", "f32"):
        f = getattr(lib, f"olrt_opd_epilogue_{suf}")
        f.restype = C.c_int
        f.argtypes = [PtrArray8, v, v, C.c_int64, C.POINTER(OlrtRefSphere), v, v, C.c_int, v]
    lib.olrt_pupil_random_f64.restype = C.c_int
    lib.olrt_pupil_random_f64.argtypes = [C.POINTER(C.c_uint64 * 2), C.POINTER(C.c_uint64 * 2), C.c_int64, v, v, v, v]
    lib.olrt_pupil_hexapolar_f64.restype = C.c_int
    lib.olrt_pupil_hexapolar_f64.argtypes = [C.c_int32, v, v, v]
    lib.olrt_pupil_uniform_f64.restype = C.c_int
    lib.olrt_pupil_uniform_f64.argtypes = [C.c_int64, v, v, v, v, v]
    lib.olrt_selftest_math_f64.restype = C.c_int
    lib.olrt_selftest_math_f64.argtypes = [v, v, v, v, C.c_int64, v]
    if lib.olrt_version() != OLRT_ABI_VERSION:
        raise OlrtLibraryError("libolrt.so ABI version mismatch; rebuild it")
    _lib = lib
    return lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise OlrtLibraryError(f"{what} failed: {ERRORS.get(rc, rc)}")
