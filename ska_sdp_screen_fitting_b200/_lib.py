"""
ctypes binding of ``libssf_b200.so`` (include/ssf.h).

There is NO CPU fallback: if the library is missing the import of any compute entry
point raises ``SsfLibraryError``.  ``check()`` turns negative status codes into
``SsfError`` carrying ``ssf_last_error()``.
"""
import ctypes
import os
from ctypes import c_char_p, c_double, c_int, c_int64, c_uint64, c_void_p

PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG, "libssf_b200.so")


class SsfLibraryError(RuntimeError):
    """libssf_b200.so is missing or cannot be loaded."""


class SsfError(RuntimeError):
    """A C-ABI call returned a negative status."""

    def __init__(self, fn, code, message):
        super().__init__(f"{fn} failed with status {code}: {message}")
        self.code = code


# name -> (restype, [argtypes]); mirrors include/ssf.h declaration by declaration
_P = c_void_p
SIGNATURES = {
    "ssf_version": (c_int, []),
    "ssf_last_error": (c_char_p, []),
    "ssf_launch_count": (c_uint64, []),
    "ssf_kl_decompose": (c_int, [_P, c_int, _P, c_int, c_double, c_double, c_double,
                                 _P, _P, _P, _P, _P]),
    "ssf_kl_patterns_update": (c_int, [_P, c_int64, c_int, _P, c_double, c_double, c_double, _P, _P, c_int,
                                       _P, c_int, _P, _P, _P, _P, _P, _P, _P, _P]),
    "ssf_kl_fit_ctas_per_sm": (c_int, [c_int]),
    "ssf_kl_fit_pass": (c_int, [_P, _P, _P, _P, c_int, c_int64, _P, _P, _P, c_double, c_double,
                                c_double, c_int, _P, _P, _P, _P, _P, _P]),
    "ssf_kl_flag_outliers": (c_int, [_P, _P, c_int, c_int64, c_double, _P, _P]),
    "ssf_kl_flag_outliers_amp": (c_int, [_P, _P, _P, c_int, c_int64, c_int64, c_double, _P, _P, _P]),
    "ssf_kl_basis": (c_int, [_P, _P, c_int, c_int, _P, c_int, _P, c_double, c_double, _P, _P]),
    "ssf_kl_coeffs": (c_int, [_P, c_int64, c_int, _P, _P, c_int, _P]),
    "ssf_kl_eval": (c_int, [_P, _P, c_int, c_int, c_int64, c_int64, _P, c_int, c_int, _P]),
    "ssf_kl_eval_tec": (c_int, [_P, _P, c_int, c_int, c_int64, c_int64, _P, c_int, c_int, _P]),
    "ssf_kl_eval_gain": (c_int, [_P, _P, c_int, c_int, c_int64, c_int64, _P, c_int, c_int, _P]),
    "ssf_voronoi_assign": (c_int, [_P, _P, c_int, c_int, c_int, _P, _P]),
    "ssf_voronoi_lut": (c_int, [_P, _P, c_int64, c_int, _P, _P]),
    "ssf_voronoi_fill": (c_int, [_P, _P, c_int, c_int64, c_int64, _P, c_int, c_int, _P]),
    "ssf_voronoi_lut_gain": (c_int, [_P, _P, c_int64, c_int, _P, _P]),
    "ssf_voronoi_fill_gain": (c_int, [_P, _P, c_int, c_int64, c_int64, _P, c_int, c_int, _P]),
    "ssf_voronoi_indicators": (c_int, [_P, c_int, c_int64, _P, _P]),
    "ssf_voronoi_contract": (c_int, [_P, _P, c_int, c_int64, c_int64, _P, c_int, c_int, _P]),
    "ssf_gauss_smooth": (c_int, [_P, c_int64, c_int, c_int, c_double, _P, _P]),
    "ssf_gauss_smooth_slab": (c_int, [_P, c_int64, c_int, c_int, c_double, _P, c_int, _P]),
    "ssf_gauss_smooth_static": (c_int, [_P, c_int64, c_int, c_int, c_double, _P]),
    "ssf_gauss_smooth_static_max_floats": (c_int64, []),
    "ssf_nan_fix": (c_int, [_P, c_int64, c_int64, _P]),
    "ssf_bswap32": (c_int, [_P, c_int64, _P]),
    "ssf_copy_rows_d2h": (c_int, [_P, c_int64, _P, c_int64, c_int64, c_int64, _P]),
    "ssf_host_place": (c_int, [_P, _P, c_int64, c_int64, c_int, c_int]),
    "ssf_voronoi_cube_host": (c_int, [_P, _P, c_int, c_int, c_int, c_int64, c_double, _P]),
    "ssf_kl_cube_host": (c_int, [_P, _P, c_int, c_int, _P, c_int, _P, c_int64, c_double, c_double,
                                 c_double, _P]),
}

OUT_BIG_ENDIAN = 1      # SSF_OUT_BIG_ENDIAN
OUT_PLANES2 = 2         # SSF_OUT_PLANES2

_loaded = {}


def load(path=None):
    """Loads ``libssf_b200.so`` (once) and declares every signature.  ``path``: another build of
    the same ABI -- the product never passes one."""
    path = path or LIB_PATH
    if path in _loaded:
        return _loaded[path]
    if not os.path.exists(path):
        raise SsfLibraryError(
            f"{path} not found: build it with `python -m ska_sdp_screen_fitting_b200._build` "
            "(there is no CPU fallback)")
    if path == LIB_PATH:
        # libcudart.so.12 is resolved through the copy torch has already mapped
        import torch  # noqa: F401  pylint: disable=import-outside-toplevel,unused-import
    try:
        lib = ctypes.CDLL(path)
    except OSError as exc:
        raise SsfLibraryError(f"cannot load {path}: {exc}") from exc
    for name, (res, args) in SIGNATURES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError as exc:
            raise SsfLibraryError(f"{path} does not export {name}") from exc
        fn.restype = res
        fn.argtypes = args
    _loaded[path] = lib
    return lib


def last_error(lib=None):
    return ((lib or load()).ssf_last_error() or b"").decode("utf8", "replace")


def launch_count(lib=None):
    return int((lib or load()).ssf_launch_count())


def call(name, *args, lib=None):
    """Calls ``name`` and raises SsfError on a negative status."""
    lib = lib or load()
    rc = getattr(lib, name)(*args)
    if rc != 0:
        raise SsfError(name, rc, last_error(lib))
    return rc
