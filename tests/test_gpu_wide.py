"""Wide GPU parity sweeps (VERDICT r1 "parity gaps"): thousands of FAST-LTS windows against the oracle, and the
FP32-screen cross-correlation against the exact time-domain kernel on adversarial records, with the screen's error
measured at EVERY lag (NBLS_OPT_SCREEN_AUDIT).

Bars: LTS outlier flags and integer lags bit-exact; velocity / sigma_tau 1e-6 relative; back-azimuth 1e-3 deg;
FP32 screen error over all searched lags < 1e-5 (half the candidate band XF_TAU32 = 2e-5, xcorr_fft.cu).
"""
import os

import numpy as np
import pytest

from narrow_band_least_squares_b200 import _capi
from narrow_band_least_squares_b200.geometry import CoArray
from narrow_band_least_squares_b200.synth import disc_geometry, plane_wave_record
from narrow_band_least_squares_b200.geometry import get_rij
from oracle import nbls_oracle as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    from narrow_band_least_squares_b200.engine import Engine
    e = Engine(0)
    yield e
    e.close()


# ---------------------------------------------------------------------------------------------
# FAST-LTS: >= 5000 windows per geometry (lts_classes.py:58-150 fast_LTS, :643-764 post_process)
# ---------------------------------------------------------------------------------------------
def synth_lags(seed, nchans, aperture_km, fs, nwin, nsamp=3000):
    """Integer lag tables the way correlate() produces them (lts_classes.py:875-876: tau = -lag/fs) for plane waves of
    random direction and speed with timing jitter, plus the outlier patterns LTS exists for: one bad element (all of
    its pairs wrong), a random fraction of wrong pairs, noise-only windows, and windows the data-spike rule skips
    (three or more zero lags, lts_classes.py:65)."""
    rng = np.random.default_rng(seed)
    lat, lon = disc_geometry(rng, nchans, aperture_km)
    rij = get_rij(lat, lon, nchans)
    co = CoArray(rij)
    xij = co.xij
    p = xij.shape[0]
    ia = np.array([i for i in range(nchans - 1) for _ in range(i + 1, nchans)])
    ib = np.array([j for i in range(nchans - 1) for j in range(i + 1, nchans)])
    lag = np.zeros((nwin, p), dtype=np.int32)
    for w in range(nwin):
        baz, vel = rng.uniform(0, 360), rng.choice([0.25, 0.34, 0.45, 3.0])
        z = np.array([np.sin(np.radians(baz)), np.cos(np.radians(baz))]) / vel
        jitter = rng.choice([0.0, 0.3, 1.0, 3.0])
        t = (xij @ z) * fs + jitter * rng.standard_normal(p)
        mode = w % 8
        if mode == 1:                                 # one bad element
            e = rng.integers(nchans)
            bad = (ia == e) | (ib == e)
            t[bad] = rng.uniform(-nsamp + 1, nsamp - 1, bad.sum())
        elif mode == 2:                               # a random 10-45 % of the pairs are wrong
            bad = rng.random(p) < rng.uniform(0.1, 0.45)
            t[bad] = rng.uniform(-nsamp + 1, nsamp - 1, bad.sum())
        elif mode == 3:                               # noise only
            t = rng.uniform(-nsamp + 1, nsamp - 1, p)
        elif mode == 4:                               # one element shifted by a constant (0.5 s in the examples)
            e = rng.integers(nchans)
            t[ia == e] += 0.5 * fs
            t[ib == e] -= 0.5 * fs
        elif mode == 5 and w % 40 == 5:               # skipped by the spike rule: >= 3 zero lags
            t[rng.choice(p, 3, replace=False)] = 0.0
        lag[w] = -np.rint(t).astype(np.int32)
    return rij, co, lag


def _oracle_lts_chunk(tau, mad, xij, xij_std, xij_mad, alpha, h):
    coeffs = orc.fast_lts(tau, mad, xij_std, xij_mad, h)
    return orc.lts_post(tau, xij, coeffs, alpha, h)


@pytest.mark.parametrize("nchans,alpha,aperture,nwin", [(6, 0.75, 0.5, 5120), (12, 0.5, 0.8, 5120), (25, 0.6, 1.0, 96)],
                         ids=["P15_a075", "P66_a05", "P300_a06"])
def test_lts_flags_bit_exact_over_5000_windows(eng, nchans, alpha, aperture, nwin):
    """P = 15: one key per lane, rank counting; P = 66: three keys per lane, rank counting; P = 300 (a
    25-element array, fewer windows: the oracle needs ~2 s per window there): ten keys per lane, the h-th key by radix select."""
    from joblib import Parallel, delayed
    fs = 100.0
    rij, co, lag = synth_lags(4000 + nchans, nchans, aperture, fs, nwin)
    p = lag.shape[1]
    cmax = np.full((nwin, p), 0.5)
    got = eng.solve(lag, cmax, rij, fs, alpha)
    tau = (-lag.T.astype(np.float64)) / fs                      # (P, W)
    mad = 1.4826 * np.median(np.abs(tau), axis=0)
    h = orc.h_value(p, alpha)
    _, xij, xij_mad, xij_std = orc.co_array(rij)
    nj = min(os.cpu_count() or 1, 16)
    edges = np.linspace(0, nwin, nj + 1).astype(int)
    parts = Parallel(n_jobs=nj)(delayed(_oracle_lts_chunk)(tau[:, a:b], mad[a:b], xij, xij_std, xij_mad, alpha, h)
                                for a, b in zip(edges[:-1], edges[1:]) if b > a)
    vel = np.concatenate([q[0] for q in parts]); baz = np.concatenate([q[1] for q in parts])
    sig = np.concatenate([q[2] for q in parts]); wts = np.concatenate([q[3] for q in parts], axis=1)
    skipped = (mad == 0) | (np.count_nonzero(tau, axis=0) < p - 2)
    assert 0 < skipped.sum() < nwin // 2
    bad = np.where(np.any(got["weights"].T != wts, axis=0))[0]
    assert bad.size == 0, (f"LTS flags differ in {bad.size} of {nwin} windows (P={p}, h={h}); first: {bad[:8]}, "
                           f"modes {bad[:8] % 8}")
    assert np.array_equal(np.isnan(got["vel"]), np.isnan(vel))
    assert np.allclose(got["vel"], vel, rtol=1e-6, atol=0, equal_nan=True)
    assert np.allclose(got["sigma_tau"], sig, rtol=1e-6, atol=1e-12, equal_nan=True)
    d = np.abs(((got["baz"] - baz + 180.0) % 360.0) - 180.0)
    assert np.nanmax(d, initial=0.0) <= 1e-3
    assert np.array_equal(got["mad"], mad)


# ---------------------------------------------------------------------------------------------
# FP32 screen under adversarial data (lts_classes.py:852-880 correlate; xcorr_fft.cu candidate band)
# ---------------------------------------------------------------------------------------------
def adversarial_record(kind, seed, nchans, npts, fs, n):
    """(C, T) 'filtered' records that stress the FP32 transform: every window of ``n`` samples sees the feature."""
    rng = np.random.default_rng(seed)
    t = np.arange(npts) / fs
    if kind == "dc_tone":            # large DC offset + a weak tone + a little noise: the energy sits in one FFT bin
        x = 50.0 + 0.05 * np.sin(2 * np.pi * 1.7 * t)[None, :] * np.ones((nchans, 1))
        x = x + 1e-3 * rng.standard_normal((nchans, npts))
        for c in range(nchans):
            x[c] = np.roll(x[c], 3 * c)
    elif kind == "spike":            # one large spike per window on every channel over quiet noise
        x = 1e-3 * rng.standard_normal((nchans, npts))
        for c in range(nchans):
            x[c, (n // 2 + 7 * c)::(n // 2)] += 100.0
    elif kind == "ramp":             # amplitude grows by 10^6 inside every window
        env = 10.0 ** (6.0 * ((np.arange(npts) % (n // 2)) / (n // 2)))
        s = rng.standard_normal(npts + 64)
        x = np.stack([s[5 * (c % 12):5 * (c % 12) + npts] for c in range(nchans)]) * env[None, :]
        x += 1e-2 * env[None, :] * rng.standard_normal((nchans, npts))
    elif kind == "sines":            # pure sines (a different phase per channel): flat-topped, periodic correlation
        f0 = 0.53
        x = np.stack([np.sin(2 * np.pi * f0 * t + 0.37 * c) for c in range(nchans)])
        x += 1e-6 * rng.standard_normal((nchans, npts))
    elif kind == "noise":            # incoherent noise: no dominant peak, many lags within a few per cent
        x = rng.standard_normal((nchans, npts))
    else:
        raise ValueError(kind)
    return np.ascontiguousarray(x)


@pytest.mark.parametrize("kind", ["dc_tone", "spike", "ramp", "sines", "noise"])
@pytest.mark.parametrize("method", [_capi.XCORR_FFT32, _capi.XCORR_AUTO], ids=["fft32", "auto"])
def test_fp32_screen_adversarial_equals_time_domain(eng, kind, method):
    """>= 10^5 pair-windows per case (10^6 over the module) at three transform lengths: the FP32-screen path returns
    the time-domain kernel's integer lags and cmax on data built to break an FP32 transform, and the audited FP32 error
    over every searched lag stays below half the candidate band."""
    fs, nchans = 100.0, 16                      # 120 pairs
    total_pw, worst = 0, 0.0
    for n in (1500, 3000, 6000):
        hop = n // 2
        nwin = 300 if n < 6000 else 240
        npts = hop * (nwin + 1) + 1
        x = adversarial_record(kind, 77, nchans, npts, fs, n)
        filt = x[None, :, :]
        ref_lag, ref_cmax = eng.xcorr(filt, [n], [hop], [nwin], method=_capi.XCORR_DIRECT)
        eng.set_option(_capi.OPT_SCREEN_AUDIT, 1)
        eng.xcorr_audit(reset=True)
        try:
            lag, cmax = eng.xcorr(filt, [n], [hop], [nwin], method=method)
        finally:
            eng.set_option(_capi.OPT_SCREEN_AUDIT, 0)
        err, nl = eng.xcorr_audit(reset=True)
        assert nl > 0, "audit did not run"
        worst = max(worst, err)
        neq = int((lag != ref_lag).sum())
        assert neq == 0, f"{kind} N={n}: {neq} of {lag.size} integer lags differ from the time-domain kernel"
        assert np.allclose(cmax, ref_cmax, rtol=0, atol=1e-11, equal_nan=True), (kind, n)
        # the lazy path (no cmax buffer: only undecided pair-windows are evaluated in FP64) must return the same lags
        lag_lazy, _ = eng.xcorr(filt, [n], [hop], [nwin], method=method, want_cmax=False)
        neq = int((lag_lazy != ref_lag).sum())
        assert neq == 0, f"{kind} N={n}: {neq} of {lag.size} lazy-path lags differ from the time-domain kernel"
        total_pw += lag.size
    assert total_pw >= 100000
    assert worst < 1e-5, f"{kind}: FP32 screen error over all lags {worst:.3g} reaches half the candidate band"


def test_screen_audit_on_plane_wave_bands(eng):
    """The audit on the realistic case: narrow-band plane-wave data of the lowest and highest c3 bands, stage-A and
    full-length transforms: error at every lag, not only at the maximum."""
    from narrow_band_least_squares_b200.configs import config_job
    j = config_job("c3")
    rec = plane_wave_record(1003, 25, 60000, 100.0, 1.0)
    worst = {}
    for b in (0, 10, 19):
        sos = [orc.design_sos("butter", j["freqlist"][b], j["freqlist"][b + 1], rec.fs, 2)]
        filt = eng.filter_bands(rec.data, rec.fs, sos)
        n = int(j["winlen_list"][b] * rec.fs)
        hop = n // 2
        nwin = len(np.arange(0, rec.npts - n, hop))
        for two_stage in (1, 0):
            eng.set_option(_capi.OPT_SCREEN_AUDIT, 1)
            eng.set_option(_capi.OPT_TWO_STAGE, two_stage)
            eng.xcorr_audit(reset=True)
            try:
                eng.xcorr(filt, [n], [hop], [nwin], method=_capi.XCORR_AUTO)
            finally:
                eng.set_option(_capi.OPT_SCREEN_AUDIT, 0)
                eng.set_option(_capi.OPT_TWO_STAGE, 1)
            err, nl = eng.xcorr_audit(reset=True)
            assert nl >= nwin * 300 * n // 2
            worst[(b, two_stage)] = err
    assert max(worst.values()) < 5e-6, worst


# ---------------------------------------------------------------------------------------------
# lazy verification (NBLS_OPT_LAZY_VERIFY): without a per-pair cmax request only the pair-windows that need it are
# evaluated in FP64 - the lags must not change by a bit, the MdCCM by no more than the last one
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name,hours,noise", [("c3", 0.25, False), ("c3", 0.2, True), ("c2", 0.5, False), ("c1", None, False)],
                         ids=["c3_plane_wave", "c3_noise", "c2_lts", "c1"])
def test_lazy_verification_changes_nothing(eng, name, hours, noise):
    from narrow_band_least_squares_b200.configs import config_plan
    from narrow_band_least_squares_b200.engine import Results
    from narrow_band_least_squares_b200.synth import config_record
    rec = config_record(name, hours=hours)
    data = rec.data
    if noise:
        data = np.random.default_rng(77).standard_normal(data.shape)      # incoherent: flat cmax distribution, many redone pairs
    plan = config_plan(name, rec.npts, rec.rij)
    out = {}
    for lazy in (1, 0):
        eng.set_option(_capi.OPT_LAZY_VERIFY, lazy)
        eng.xcorr_stats(reset=True)
        try:
            res = eng.run(plan, data, Results(plan, pinned=False, want_cmax=False))
        finally:
            eng.set_option(_capi.OPT_LAZY_VERIFY, 1)
        st = eng.xcorr_stats(reset=True)
        out[lazy] = (res, st)
    full = eng.run(plan, data, Results(plan, pinned=False, want_cmax=True))      # a cmax request: dense evaluation, exact cmax
    a, b = out[1][0], out[0][0]
    assert np.array_equal(a.lag, b.lag) and np.array_equal(a.lag, full.lag)
    for k in ("mad", "vel", "baz", "sigma_tau", "slowness"):
        assert np.array_equal(getattr(a, k), getattr(b, k), equal_nan=True), k
        assert np.array_equal(getattr(a, k), getattr(full, k), equal_nan=True), k
    # MdCCM: the same middle pairs, evaluated exactly; the item kernel and the dense kernel add the same FP64 products in a
    # different order, hence the last bit (bar for MdCCM: 1e-9)
    assert np.array_equal(np.isnan(a.mdccm), np.isnan(b.mdccm))
    assert np.nanmax(np.abs(a.mdccm - b.mdccm), initial=0.0) <= 1e-15
    assert np.array_equal(b.mdccm, full.mdccm, equal_nan=True)
    if a.weights is not None:
        assert np.array_equal(a.weights, b.weights)
    # the MdCCM really is the median of the exact per-pair maxima
    assert np.allclose(a.mdccm, np.nanmedian(full.cmax, axis=1), rtol=0, atol=1e-15, equal_nan=True)
    # and the lazy run evaluated far fewer candidates exactly
    va, vb = out[1][1]["verified"], out[0][1]["verified"]
    assert 0 < va < vb, (va, vb)
    if not noise and name == "c3":
        assert va < 0.35 * vb, (va, vb)


# ---------------------------------------------------------------------------------------------
# config 5: a batch of independent arrays, OLS + LTS sweep (batch.py)
# ---------------------------------------------------------------------------------------------
def test_batch_equals_standalone_runs(eng):
    """Every array of a batch returns, for every alpha of the sweep, exactly the 9-tuple of a stand-alone driver call
    (the LTS leg re-solves on the lags the OLS leg's cross-correlation left on the device)."""
    from narrow_band_least_squares_b200 import narrow_band_least_squares, narrow_band_least_squares_batch
    from narrow_band_least_squares_b200.engine import Engine
    fs, nb = 50.0, 3
    fl = orc.get_freqlist_log(0.5, 5.0, nb)
    wl = [20.0, 16.0, 12.0]
    frl = np.logspace(-2, np.log10(fs / 2), 20)
    recs = [plane_wave_record(seed=300 + a, nchans=5 + a % 2, npts=9000 + 500 * a, fs=fs, aperture_km=0.4,
                              flip_channel=2 if a % 2 else None) for a in range(5)]
    records = [((r.data, r.fs, None), None, None, r.rij) for r in recs]
    common = (nb, np.zeros(20), np.zeros(20, complex), fl, "log", frl, "butter", 2, 0.01)
    engines = [Engine(0), Engine(0)]
    try:
        got = narrow_band_least_squares_batch(records, wl, 0.5, 1.0, *common, alphas=(1.0, 0.75), engines=engines)
    finally:
        for e in engines:
            e.close()
    assert [a for a, _ in got] == list(range(5))
    for a, per_alpha in got:
        for alpha in (1.0, 0.75):
            ref = narrow_band_least_squares(wl, 0.5, alpha, records[a][0], None, None, *common, rij=recs[a].rij, engine=eng)
            out = per_alpha[alpha]
            for k in (0, 1, 2, 3, 5):
                assert np.array_equal(out[k], ref[k], equal_nan=True), (a, alpha, k)
            assert out[6] == ref[6]
            if alpha == 1.0:
                assert out[4] is None and ref[4] is None
            else:
                assert set(out[4]) == set(ref[4]) and 'size' in out[4]
                for key in ref[4]:
                    assert np.array_equal(out[4][key], ref[4][key]), (a, key)


# ---------------------------------------------------------------------------------------------
# long windows: any window length is served; beyond the FFT plans the time-domain kernels take over
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n,max_lag", [(8192, -1), (13000, -1), (20000, -1), (20000, 700)])
def test_long_windows_match_oracle(eng, n, max_lag):
    """N = 8192 (beyond the 12288-point FFT plans: shared-memory time-domain kernel), N = 13000 and 20000 (200 s at
    100 sps, the reference example's FMIN = 0.1 Hz band: tiled time-domain kernel) against np.correlate + argmax."""
    fs = 100.0
    rec = plane_wave_record(seed=50 + n % 7, nchans=4, npts=n * 2 + 500, fs=fs, aperture_km=0.6, noise_every_min=0)
    sos = [orc.design_sos("butter", 0.1, 0.2, fs, 2)]
    filt = eng.filter_bands(rec.data, fs, sos)
    filt[0, 2, :] = 0.0 if n == 13000 else filt[0, 2, :]          # a dead channel on one case
    hop = n // 2
    starts = np.arange(0, rec.npts - n, hop)
    pairs, _, _, _ = orc.co_array(rec.rij)
    lag, cmax, *_ = orc.correlate_windows(np.ascontiguousarray(filt[0].T), starts, n, fs, pairs,
                                          max_lag=None if max_lag < 0 else max_lag)
    for method in (_capi.XCORR_AUTO, _capi.XCORR_DIRECT):
        got_lag, got_cmax = eng.xcorr(filt, [n], [hop], [len(starts)], max_lag=max_lag, method=method)
        assert np.array_equal(got_lag.T, lag), (n, method)
        assert np.allclose(got_cmax.T, cmax, rtol=0, atol=1e-12, equal_nan=True), (n, method)


# ---------------------------------------------------------------------------------------------
# tensor-pipe experiment (tc_dft.cu): the numbers of both kernels are right before their times are compared
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("mode,tol", [(0, 2e-6), (2, 2e-6), (1, 3e-3)], ids=["scalar", "tc_split3", "tc_single"])
def test_experiment_dft_pass_numbers(eng, mode, tol):
    """One radix-32 pass (32-point inverse DFT per row + the inter-pass twiddle) against float64, for the scalar
    butterflies, the three-product split-FP16 tcgen05 path (FP32-grade) and the single-product path (FP16-grade)."""
    rng = np.random.default_rng(9)
    rows = 128 * 37 + 5                       # ragged last tile
    x = (rng.standard_normal((rows, 32)) + 1j * rng.standard_normal((rows, 32))).astype(np.complex64)
    y, ms = eng.experiment_dft_pass(x, mode, reps=1, iters=1)
    n = np.arange(32)
    dft = np.exp(2j * np.pi * np.outer(n, n) / 32)                         # inverse sign, unnormalised
    w1 = np.exp(2j * np.pi * (np.arange(rows) % 256) / 8192)
    ref = (x.astype(np.complex128) @ dft.T) * (w1[:, None] ** n[None, :])
    err = np.max(np.abs(y - ref)) / np.max(np.abs(ref))
    assert err < tol, (mode, err)
    assert ms > 0


# ---------------------------------------------------------------------------------------------
# filter: higher orders on a narrow band (the cascade's transition matrix is non-normal there)
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("order", [2, 4, 8])
def test_filter_high_order_narrow_band(eng, order):
    """Butterworth orders 2 / 4 / 8 on the narrowest c3 band (0.5-0.561 Hz at 100 sps: poles within 0.002 of each other
    and of the unit circle).  Judge: a LONG-DOUBLE sosfilt; the multi-chunk scan with truncated state passes must be as
    close to it as scipy's own float64 sosfilt is (both carry ~1e-12 of the RMS in rounding at this conditioning)."""
    from scipy import signal
    fs = 100.0
    rng = np.random.default_rng(order)
    x = rng.standard_normal((2, 600000))
    sos = orc.design_sos("butter", 0.5, 0.5610092271509818, fs, order)
    wlen, up, down = orc.taper_weights(x.shape[1], 0.01)
    exact = np.empty(x.shape, dtype=np.longdouble)
    f64 = np.empty(x.shape)
    for c in range(2):
        y = signal.sosfilt(sos.astype(np.longdouble), x[c].astype(np.longdouble))
        exact[c] = signal.sosfilt(sos.astype(np.longdouble), y[::-1])[::-1]
        y = signal.sosfilt(sos, x[c])
        f64[c] = signal.sosfilt(sos, y[::-1])[::-1]
    for arr in (exact, f64):
        arr[:, :wlen] *= up
        arr[:, x.shape[1] - wlen:] *= down
    rms = np.sqrt(np.mean(np.asarray(exact, dtype=np.float64) ** 2, axis=1, keepdims=True))
    err_scipy = float(np.max(np.abs(f64 - exact) / rms))
    try:
        for sub in (1, 4):          # forward state pass on whole chunks (default) and on sub-chunks (NBLS_OPT_FILTER_P1_SUB)
            eng.set_option(_capi.OPT_FILTER_P1_SUB, sub)
            for kw in (dict(chunk_len=8192), dict(chunk_len=65536), dict(chunk_len=8192, exact_scan=True), dict()):
                out = eng.filter_bands(x, fs, [sos], zerophase=True, **kw)[0]
                err = float(np.max(np.abs(out - exact) / rms))
                assert err <= max(3.0 * err_scipy, 3e-12), (order, sub, kw, err, err_scipy)
    finally:
        eng.set_option(_capi.OPT_FILTER_P1_SUB, 1)
