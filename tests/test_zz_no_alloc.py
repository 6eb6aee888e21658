"""nbls_xcorr_static: the allocation-free cross-correlation entry point (scratch in a static __device__ array of the
library; include/nbls.h) against the existing nbls_xcorr entry point, through the C ABI, on two calls in a row.

Kept in a file of its own that pytest collects LAST.  Passes alone in a fresh process
and after the other GPU tests: it creates its own context and shares no state with them.
"""
import os

import numpy as np
import pytest

from narrow_band_least_squares_b200 import _capi
from narrow_band_least_squares_b200.engine import DeviceArray, Engine

pytestmark = pytest.mark.gpu

CMAX_TOL = 1e-12      # absolute, the bar of the other cross-correlation tests


def _record(seed, c, t):
    """Band-limited noise plus a common delayed wavelet: distinct, well-separated correlation maxima."""
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((c, t))
    k = np.hanning(25)
    x = np.stack([np.convolve(r, k, mode="same") for r in x])
    common = np.convolve(rng.standard_normal(t + 64), k, mode="same")
    for i in range(c):
        x[i] += 2.0 * common[3 * i:3 * i + t]
    return x


def _static(eng, filt, n, hop, first, w, max_lag, want_cmax=True):
    c, t = filt.shape
    ld = (t + 15) // 16 * 16
    ypad = np.zeros((c, ld))
    ypad[:, :t] = filt
    p = c * (c - 1) // 2
    yd = DeviceArray(eng, ypad.nbytes).upload(ypad)
    lag_d = DeviceArray(eng, 4 * w * p)
    cm_d = DeviceArray(eng, 8 * w * p)
    try:
        # poison the outputs so that a kernel that never ran cannot pass
        lag_d.upload(np.full((w, p), 123456, dtype=np.int32))
        cm_d.upload(np.full((w, p), -7.0))
        _capi.check(eng.lib.nbls_xcorr_static(eng.ctx, yd.ptr, c, ld, n, hop, first, w, max_lag, lag_d.ptr,
                                              cm_d.ptr if want_cmax else None))
        eng.synchronize()
        return lag_d.download((w, p), np.int32), cm_d.download((w, p), np.float64)
    finally:
        yd.free(); lag_d.free(); cm_d.free()


def test_static_scratch_xcorr_matches_nbls_xcorr_twice():
    eng = Engine(0)
    # call 1: a short window (nbls_xcorr runs the one-CTA-per-pair kernel: a different summation order, so cmax to tolerance);
    # call 2: other data, a window long enough for several lag blocks and a restricted lag range; then call 1 again, so stale
    # scratch contents of a LARGER call sit under a smaller one
    cases = [
        dict(seed=1, c=5, t=6000, n=1500, hop=750, first=40, max_lag=-1),
        dict(seed=2, c=4, t=21000, n=5000, hop=4000, first=0, max_lag=-1),
        dict(seed=3, c=4, t=9000, n=3000, hop=1500, first=7, max_lag=700),
        dict(seed=1, c=5, t=6000, n=1500, hop=750, first=40, max_lag=-1),
    ]
    results = []
    for k in cases:
        x = _record(k["seed"], k["c"], k["t"])
        w = (k["t"] - k["first"] - k["n"]) // k["hop"] + 1
        lag_ref, cmax_ref = eng.xcorr(x[None], [k["n"]], [k["hop"]], [w], win_first=[k["first"]], max_lag=k["max_lag"],
                                      method=_capi.XCORR_DIRECT)
        lag, cmax = _static(eng, x, k["n"], k["hop"], k["first"], w, k["max_lag"])
        assert lag.shape == lag_ref.shape and w >= 2
        np.testing.assert_array_equal(lag, lag_ref)
        assert np.max(np.abs(cmax - cmax_ref)) <= CMAX_TOL
        assert np.all(np.abs(lag) <= (k["n"] - 1 if k["max_lag"] < 0 else k["max_lag"]))
        results.append((lag, cmax))
    # the repeated case returns bit-identical results whatever ran in between
    np.testing.assert_array_equal(results[0][0], results[3][0])
    np.testing.assert_array_equal(results[0][1], results[3][1])


def test_static_scratch_xcorr_dead_channel_and_null_cmax():
    eng = Engine(0)
    x = _record(5, 4, 5000)
    x[2] = 0.0                                   # zero-energy channel: lag -(N-1), cmax NaN (lts_classes.py:871-875)
    n, hop = 1000, 500
    w = (5000 - n) // hop + 1
    lag_ref, cmax_ref = eng.xcorr(x[None], [n], [hop], [w], method=_capi.XCORR_DIRECT)
    lag, cmax = _static(eng, x, n, hop, 0, w, -1)
    np.testing.assert_array_equal(lag, lag_ref)
    np.testing.assert_array_equal(np.isnan(cmax), np.isnan(cmax_ref))
    assert np.isnan(cmax).any()
    ok = ~np.isnan(cmax)
    assert np.max(np.abs(cmax[ok] - cmax_ref[ok])) <= CMAX_TOL
    lag2, cm2 = _static(eng, x, n, hop, 0, w, -1, want_cmax=False)
    np.testing.assert_array_equal(lag2, lag_ref)
    assert np.all(cm2 == -7.0)                   # cmax_dev = NULL: nothing written


def test_static_scratch_xcorr_rejects_oversized_and_bad_requests():
    eng = Engine(0)
    lib = eng.lib
    header = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "nbls.h")
    assert "#define NBLS_XCORR_STATIC_MAX_ENTRIES 262144" in open(header).read()
    dummy = DeviceArray(eng, 4096)
    try:
        # 25 elements (300 pairs), 30 s windows at 100 sps (3 lag blocks), 300 windows: 270 000 entries > 262 144
        launches = lib.nbls_launch_count(eng.ctx)
        rc = lib.nbls_xcorr_static(eng.ctx, dummy.ptr, 25, 10 ** 7, 3000, 1500, 0, 300, -1, dummy.ptr, None)
        assert rc != 0
        msg = lib.nbls_last_error().decode()
        assert "NBLS_XCORR_STATIC_MAX_ENTRIES" in msg and "270000" in msg
        assert lib.nbls_launch_count(eng.ctx) == launches      # nothing was launched
        # windows running past the row, null pointers, one channel
        assert lib.nbls_xcorr_static(eng.ctx, dummy.ptr, 4, 1000, 600, 300, 0, 3, -1, dummy.ptr, None) != 0
        assert "ldy" in lib.nbls_last_error().decode()
        assert lib.nbls_xcorr_static(eng.ctx, None, 4, 1000, 100, 50, 0, 3, -1, dummy.ptr, None) != 0
        assert lib.nbls_xcorr_static(eng.ctx, dummy.ptr, 1, 1000, 100, 50, 0, 3, -1, dummy.ptr, None) != 0
        assert lib.nbls_xcorr_static(None, dummy.ptr, 4, 1000, 100, 50, 0, 3, -1, dummy.ptr, None) != 0
        # an empty request is fine and launches nothing
        assert lib.nbls_xcorr_static(eng.ctx, dummy.ptr, 4, 1000, 100, 50, 0, 0, -1, dummy.ptr, None) == 0
        assert lib.nbls_launch_count(eng.ctx) == launches
    finally:
        dummy.free()
