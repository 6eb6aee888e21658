"""ctypes binding of libnbls.so (include/nbls.h) — the only route from Python to the CUDA kernels.

There is no CPU fallback: if the shared library has not been built, or no CUDA device is
present, the first call raises.  Build with ``python -c "import __graft_entry__ as g; g.build()"``
or ``make -C narrow_band_least_squares_b200/csrc``.
"""
from __future__ import annotations

import ctypes as C
import os
import weakref

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_lib", "libnbls.so")

c_double_p = C.POINTER(C.c_double)
c_int32_p = C.POINTER(C.c_int32)
c_int64_p = C.POINTER(C.c_int64)
c_uint8_p = C.POINTER(C.c_uint8)

XCORR_AUTO, XCORR_DIRECT, XCORR_FFT, XCORR_FFT32 = 0, 1, 2, 3
# nbls_set_option keys (include/nbls.h)
(OPT_PROFILE, OPT_TWO_STAGE, OPT_FILTER_SPLIT, OPT_FILTER_SHARED_P1, OPT_FILTER_LANES, OPT_LTS_FP64_RANK,
 OPT_SCREEN_AUDIT, OPT_FILTER_P1_WARPS, OPT_FILTER_P1_SUB, OPT_FILTER_OVERLAP, OPT_LAZY_VERIFY) = range(11)


class FilterDesc(C.Structure):
    _fields_ = [("nbands", C.c_int32), ("nsections", C.c_int32), ("sos", c_double_p), ("zerophase", C.c_int32),
                ("exact_scan", C.c_int32), ("chunk_len", C.c_int32), ("reserved0", C.c_int32),
                ("taper_len", C.c_int64), ("taper_up", c_double_p), ("taper_down", c_double_p),
                ("global_t0", C.c_int64), ("global_npts", C.c_int64)]


class WindowDesc(C.Structure):
    _fields_ = [("nbands", C.c_int32), ("max_lag", C.c_int32), ("method", C.c_int32), ("reserved0", C.c_int32),
                ("win_len", c_int32_p), ("win_hop", c_int32_p), ("win_first", c_int64_p), ("win_count", c_int32_p)]


class SolveDesc(C.Structure):
    _fields_ = [("nchan", C.c_int32), ("npairs", C.c_int32), ("fs", C.c_double), ("xij", c_double_p),
                ("xij_mad", c_double_p), ("ols_gain", c_double_p), ("alpha", C.c_double), ("h", C.c_int32),
                ("n_starts", C.c_int32), ("n_keep", C.c_int32), ("csteps1", C.c_int32), ("csteps2", C.c_int32),
                ("reserved0", C.c_int32), ("subsets", c_int32_p), ("raw_factor", C.c_double),
                ("rew_factor1", C.c_double), ("rew_cons", c_double_p)]


class Outputs(C.Structure):
    _fields_ = [("mdccm", C.c_void_p), ("mad", C.c_void_p), ("vel", C.c_void_p), ("baz", C.c_void_p),
                ("sigma_tau", C.c_void_p), ("slowness", C.c_void_p), ("lag", C.c_void_p), ("cmax", C.c_void_p),
                ("weights", C.c_void_p)]


class Job(C.Structure):
    _fields_ = [("filter", FilterDesc), ("window", WindowDesc), ("solve", SolveDesc)]


EXPORTS = {
    # name: (restype, argtypes)
    "nbls_version": (C.c_int, []),
    "nbls_last_error": (C.c_char_p, []),
    "nbls_device_count": (C.c_int, [c_int32_p]),
    "nbls_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_int32]),
    "nbls_destroy": (C.c_int, [C.c_void_p]),
    "nbls_synchronize": (C.c_int, [C.c_void_p]),
    "nbls_set_workspace_limit": (C.c_int, [C.c_void_p, C.c_size_t]),
    "nbls_malloc_device": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.c_size_t]),
    "nbls_free_device": (C.c_int, [C.c_void_p, C.c_void_p]),
    "nbls_malloc_host": (C.c_int, [C.POINTER(C.c_void_p), C.c_size_t]),
    "nbls_free_host": (C.c_int, [C.c_void_p]),
    "nbls_host_register": (C.c_int, [C.c_void_p, C.c_size_t]),
    "nbls_host_unregister": (C.c_int, [C.c_void_p]),
    "nbls_memcpy_h2d": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]),
    "nbls_memcpy_d2h": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]),
    "nbls_filter": (C.c_int, [C.c_void_p, C.POINTER(FilterDesc), C.c_void_p, C.c_int32, C.c_int64, C.c_int64,
                              C.c_void_p, C.c_int64, C.c_int64]),
    "nbls_xcorr": (C.c_int, [C.c_void_p, C.POINTER(WindowDesc), C.c_void_p, C.c_int32, C.c_int64, C.c_int64,
                             C.c_void_p, C.c_void_p]),
    "nbls_xcorr_fft_length": (C.c_int, [C.c_int32, C.c_int32]),
    "nbls_xcorr_stats": (C.c_int, [C.c_void_p, c_int64_p, c_int64_p, c_double_p, C.c_int32]),
    "nbls_xcorr_two_stage_stats": (C.c_int, [C.c_void_p, c_int64_p, c_int64_p, C.c_int32]),
    "nbls_xcorr_audit": (C.c_int, [C.c_void_p, c_double_p, c_int64_p, C.c_int32]),
    "nbls_xcorr_two_stage_pairs": (C.c_int, [C.c_void_p, c_int64_p, C.c_int32]),
    "nbls_set_option": (C.c_int, [C.c_void_p, C.c_int32, C.c_int64]),
    "nbls_get_option": (C.c_int, [C.c_void_p, C.c_int32, c_int64_p]),
    "nbls_kernel_group_count": (C.c_int, []),
    "nbls_kernel_group_name": (C.c_char_p, [C.c_int32]),
    "nbls_profile": (C.c_int, [C.c_void_p, C.c_int32, c_double_p, c_int64_p, c_double_p, C.c_int32]),
    "nbls_solve": (C.c_int, [C.c_void_p, C.POINTER(SolveDesc), C.c_int64, C.c_void_p, C.c_void_p,
                             C.POINTER(Outputs)]),
    "nbls_run_device": (C.c_int, [C.c_void_p, C.POINTER(Job), C.c_void_p, C.c_int32, C.c_int64, C.c_int64,
                                  C.POINTER(Outputs)]),
    "nbls_run_host": (C.c_int, [C.c_void_p, C.POINTER(Job), C.c_void_p, C.c_int32, C.c_int64, C.c_int64,
                                C.POINTER(Outputs)]),
    "nbls_resolve_host": (C.c_int, [C.c_void_p, C.POINTER(SolveDesc), C.POINTER(Outputs)]),
    "nbls_last_timings": (C.c_int, [C.c_void_p, C.POINTER(C.c_float), C.c_int32]),
    "nbls_mark": (C.c_int, [C.c_void_p, C.c_int32]),
    "nbls_mark_elapsed": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.POINTER(C.c_float)]),
    "nbls_launch_count": (C.c_int64, [C.c_void_p]),
    "nbls_experiment_dft_pass": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_void_p,
                                           C.POINTER(C.c_float)]),
    "nbls_xcorr_static": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int64, C.c_int32, C.c_int32, C.c_int64, C.c_int32,
                                    C.c_int32, C.c_void_p, C.c_void_p]),
    "nbls_measure_fp64_peak": (C.c_int, [C.c_void_p, C.c_int32, c_double_p]),
    "nbls_measure_fp32_peak": (C.c_int, [C.c_void_p, C.c_int32, c_double_p]),
}

_lib = None


class NBLSError(RuntimeError):
    pass


def load():
    """Load libnbls.so and declare every prototype.  Raises if the library is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise NBLSError(f"{LIB_PATH} not found: build the CUDA extension first (no CPU fallback exists)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in EXPORTS.items():
        fn = getattr(lib, name)     # AttributeError if the symbol is missing
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        raise NBLSError(f"libnbls error {rc}: {load().nbls_last_error().decode(errors='replace')}")


def dptr(a, typ=c_double_p):
    """Pointer to a C-contiguous numpy array (the caller keeps the array alive)."""
    return a.ctypes.data_as(typ)


def pinned_empty(shape, dtype=np.float64):
    """A numpy array over page-locked host memory (freed when the array is collected)."""
    lib = load()
    dtype = np.dtype(dtype)
    n = int(np.prod(shape)) * dtype.itemsize
    p = C.c_void_p()
    check(lib.nbls_malloc_host(C.byref(p), max(n, 1)))
    buf = (C.c_char * max(n, 1)).from_address(p.value)
    weakref.finalize(buf, _free_pinned, p.value)      # buf outlives every numpy view of it
    return np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)


def _free_pinned(addr):
    if _lib is not None:
        _lib.nbls_free_host(C.c_void_p(addr))
