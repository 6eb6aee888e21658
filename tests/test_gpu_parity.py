"""GPU parity: the CUDA path (through the C ABI of libnbls) against the golden fixtures produced by the
UNMODIFIED reference (tests/golden, oracle/run_reference.py) and against the CPU oracle.

Bars (BASELINE.json north_star): integer lags and LTS outlier flags bit-exact; trace velocity and
sigma_tau 1e-6 relative; back-azimuth 1e-3 deg; mdccm 1e-9 absolute; NaN pattern identical.
The filtered samples are compared at 1e-12 of the channel RMS (scipy restatement, parity unpinned
upstream — SURVEY.md §8(c)).
"""
import os

import numpy as np
import pytest

from narrow_band_least_squares_b200 import _capi
from narrow_band_least_squares_b200.synth import plane_wave_record
from oracle import nbls_oracle as orc
from oracle.run_reference import CASES

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture(scope="module")
def eng():
    from narrow_band_least_squares_b200.engine import Engine
    e = Engine(0)
    yield e
    e.close()


def _load(name):
    return np.load(os.path.join(GOLD, name + ".npz"), allow_pickle=False)


def _case(name):
    g = _load(name)
    skw, pkw = CASES[name]
    rec = plane_wave_record(**skw)
    tvec = (rec.starttime + np.arange(rec.npts) / rec.fs) / 86400.0
    return g, rec, tvec, pkw


def _close_est(got, ref, what):
    assert np.array_equal(np.isnan(got), np.isnan(ref)), f"{what}: NaN pattern differs"
    if what == "baz":
        d = np.abs(((got - ref + 180.0) % 360.0) - 180.0)
        assert np.nanmax(d, initial=0.0) <= 1e-3, f"baz differs by {np.nanmax(d)} deg"
    elif what == "mdccm":
        assert np.allclose(got, ref, rtol=0, atol=1e-9, equal_nan=True), what
    else:
        assert np.allclose(got, ref, rtol=1e-6, atol=1e-12, equal_nan=True), what


@pytest.mark.parametrize("name", sorted(CASES))
def test_filter_matches_oracle(eng, name):
    g, rec, _, pkw = _case(name)
    fl = g["freqlist"]
    nb = int(g["nbands"])
    sos = [orc.design_sos("butter", fl[b], fl[b + 1], rec.fs, int(g["order"])) for b in range(nb)]
    for kw in (dict(), dict(chunk_len=256, exact_scan=True), dict(chunk_len=1 << 20)):
        out = eng.filter_bands(rec.data, rec.fs, sos, zerophase=True, **kw)
        for b in range(nb):
            ref, _ = orc.filter_record(rec.data, rec.fs, "butter", fl[b], fl[b + 1], int(g["order"]))
            rms = np.sqrt(np.mean(ref ** 2, axis=1, keepdims=True))
            rms[rms == 0] = 1.0
            err = np.max(np.abs(out[b] - ref) / rms)
            assert err < 1e-12, (name, b, kw, err)
            if kw.get("chunk_len", 0) >= rec.npts:       # one chunk = scipy's sequential order: bit-exact
                assert np.array_equal(out[b], ref), (name, b, "single-chunk run must be bit-identical")
            # fixture cross-check (head/tail samples written by the reference run)
            assert np.allclose(out[b][:, :64], g[f"b{b}_filt_head"], rtol=0, atol=1e-12 * float(rms.max()))


@pytest.mark.parametrize("method", [_capi.XCORR_DIRECT, _capi.XCORR_FFT, _capi.XCORR_FFT32],
                         ids=["direct", "fft64", "fft32"])
@pytest.mark.parametrize("name", sorted(CASES))
def test_xcorr_lags_bit_exact(eng, name, method):
    g, rec, _, pkw = _case(name)
    fl, wl = g["freqlist"], g["winlen_list"]
    nb = int(g["nbands"])
    filt = np.stack([orc.filter_record(rec.data, rec.fs, "butter", fl[b], fl[b + 1], int(g["order"]))[0]
                     for b in range(nb)])
    n = [int(g[f"b{b}_winlensamp"]) for b in range(nb)]
    hop = [int(g[f"b{b}_sampinc"]) for b in range(nb)]
    cnt = [len(g[f"b{b}_intervals"]) for b in range(nb)]
    lag, cmax = eng.xcorr(filt, n, hop, cnt, method=method)
    row = 0
    for b in range(nb):
        r = slice(row, row + cnt[b])
        assert np.array_equal(lag[r].T, g[f"b{b}_lag"]), (name, b, "integer lags must match the reference exactly")
        assert np.allclose(cmax[r].T, g[f"b{b}_cmax"], rtol=0, atol=1e-12, equal_nan=True)
        row += cnt[b]


@pytest.mark.parametrize("name", sorted(CASES))
def test_solve_matches_reference(eng, name):
    g, rec, _, pkw = _case(name)
    alpha = float(g["alpha"])
    for b in range(int(g["nbands"])):
        lag = np.ascontiguousarray(g[f"b{b}_lag"].T)
        cmax = np.ascontiguousarray(g[f"b{b}_cmax"].T)
        out = eng.solve(lag, cmax, g["b0_rij"], rec.fs, alpha)
        _close_est(out["mdccm"], g[f"b{b}_mdccm"], "mdccm")
        assert np.array_equal(out["mad"], g[f"b{b}_mad"])
        _close_est(out["vel"], g[f"b{b}_vel"], "vel")
        _close_est(out["baz"], g[f"b{b}_baz"], "baz")
        _close_est(out["sigma_tau"], g[f"b{b}_sigma_tau"], "sigma_tau")
        if alpha != 1.0:
            assert np.array_equal(out["weights"].T, g[f"b{b}_element_weights"]), (name, b, "LTS flags must match exactly")


@pytest.mark.parametrize("name", sorted(CASES))
def test_driver_matches_reference_driver(eng, name):
    from narrow_band_least_squares_b200 import narrow_band_least_squares
    g, rec, tvec, pkw = _case(name)
    nb = int(g["nbands"])
    frl = g["nbls_freq_resp_list"]
    out = narrow_band_least_squares(list(g["winlen_list"]), float(g["winover"]), float(g["alpha"]),
                                    (rec.data, rec.fs, tvec), rec.lat, rec.lon, nb, np.zeros(50), np.zeros(50, complex),
                                    g["freqlist"], "log", frl, "butter", int(g["order"]), 0.01, rij=g["b0_rij"], engine=eng,
                                    return_details=True)
    (vel, baz, md, tt, stdict_all, sig, ncl, w_arr, h_arr), plan, res = out
    assert ncl == g["nbls_num_compute"].tolist()
    assert vel.shape == g["nbls_vel"].shape
    _close_est(vel, g["nbls_vel"], "vel")
    _close_est(baz, g["nbls_baz"], "baz")
    _close_est(md, g["nbls_mdccm"], "mdccm")
    assert np.array_equal(tt, g["nbls_t"], equal_nan=True)
    assert np.array_equal(w_arr, g["nbls_w"]) and np.array_equal(h_arr, g["nbls_h"])
    for b in range(nb):       # end-to-end lags (GPU-filtered input) against the reference's
        assert np.array_equal(res.band(b)["lag"], g[f"b{b}_lag"]), (name, b)
    if float(g["alpha"]) == 1.0:
        _close_est(sig, g["nbls_sig_tau"], "sigma_tau")
        assert stdict_all is None
    else:
        keys = [k for k in stdict_all if k != "size"]
        assert keys == [str(k) for k in g["nbls_stdict_keys"]]
        off = g["nbls_stdict_off"]
        for i, k in enumerate(keys):
            assert np.array_equal(stdict_all[k], g["nbls_stdict_vals"][off[i]:off[i + 1]])
        assert stdict_all["size"] == int(g["nbls_stdict_size"])


def test_ltsva_surface_and_restricted_lag(eng):
    from narrow_band_least_squares_b200 import ltsva
    g, rec, tvec, pkw = _case("g2_lts_c6")
    fl = g["freqlist"]
    filt, _ = orc.filter_record(rec.data, rec.fs, "butter", fl[0], fl[1], int(g["order"]))
    vel, baz, t, md, stdict, sig, civ, cib = ltsva((filt, rec.fs, tvec), rec.lat, rec.lon, float(g["winlen_list"][0]),
                                                   float(g["winover"]), float(g["alpha"]), rij=g["b0_rij"], engine=eng)
    _close_est(vel, g["b0_vel"], "vel")
    _close_est(baz, g["b0_baz"], "baz")
    assert np.array_equal(t, g["b0_t"])
    keys = [k for k in stdict if k != "size"]
    assert keys == [str(k) for k in g["b0_stdict_keys"]] and stdict["size"] == rec.nchans
    # slowness-ellipse uncertainties: where the origin lies outside the confidence ellipse (every usable estimate) they
    # agree with the reference's quartic-root machinery; inside it the reference's numbers are the real parts of
    # complex roots and this implementation reports inf / 180 deg (uncertainty.py)
    rv, rb = g["b0_conf_int_vel"], g["b0_conf_int_baz"]
    m = np.isfinite(rv) & np.isfinite(civ)
    assert m.sum() >= 0.5 * np.isfinite(rv).sum()
    assert np.allclose(civ[m], rv[m], rtol=1e-5, atol=0)
    assert np.allclose(cib[m], rb[m], rtol=0, atol=1e-6)
    assert np.array_equal(np.isnan(civ), np.isnan(vel))
    # restricted lag search against the oracle's sliced arg-max
    L = 150
    n, hop, starts = orc.window_grid(rec.npts, rec.fs, float(g["winlen_list"][0]), float(g["winover"]))
    pairs, *_ = orc.co_array(g["b0_rij"])
    ref_lag, ref_cmax, *_ = orc.correlate_windows(np.ascontiguousarray(filt.T), starts, n, rec.fs, pairs, None, L)
    for method in (_capi.XCORR_DIRECT, _capi.XCORR_FFT, _capi.XCORR_FFT32):
        lag, cmax = eng.xcorr(filt[None], [n], [hop], [len(starts)], max_lag=L, method=method)
        assert np.array_equal(lag.T, ref_lag), method
        assert np.allclose(cmax.T, ref_cmax, rtol=0, atol=1e-12, equal_nan=True)


def _oracle_xcorr(x, n, hop, count, fs=100.0, max_lag=None):
    c = x.shape[0]
    rij = np.vstack([np.arange(c, dtype=float), np.arange(c, dtype=float) ** 2])
    pairs, *_ = orc.co_array(rij)
    starts = np.arange(count) * hop
    lag, cmax, *_ = orc.correlate_windows(np.ascontiguousarray(x.T), starts, n, fs, pairs, None, max_lag)
    return lag.T, cmax.T


def test_xcorr_fft_ties_and_unequal_amplitudes(eng):
    """Exact ties (two equal peaks: np.argmax keeps the first) are settled by the time-domain kernel; channels of very
    different amplitude share one packed transform without losing the weak one's arg-max; dead channel -> NaN."""
    rng = np.random.default_rng(5)
    n, hop, count = 700, 350, 6
    t = n + hop * (count - 1) + 3
    x = np.zeros((5, t))
    tt = np.arange(t) % hop               # spikes repeating with the hop: every window sees the same picture
    x[0] = ((tt == 100) | (tt == 300)).astype(float)   # c_01 has two exactly equal maxima, lags 0 and +200
    x[1] = (tt == 100).astype(float)
    x[2] = rng.standard_normal(t) * 1e9   # strong rider
    x[3] = np.roll(x[2], 7) * 1e-15 + rng.standard_normal(t) * 1e-16   # weak rider, delayed copy
    x[4] = 0.0                            # dead channel
    ref_lag, ref_cmax = _oracle_xcorr(x, n, hop, count)
    for method in (_capi.XCORR_DIRECT, _capi.XCORR_FFT, _capi.XCORR_FFT32):
        lag, cmax = eng.xcorr(x[None], [n], [hop], [count], method=method)
        assert np.array_equal(lag, ref_lag), method
        assert np.allclose(cmax, ref_cmax, rtol=0, atol=1e-12, equal_nan=True), method


@pytest.mark.parametrize("method", [_capi.XCORR_FFT, _capi.XCORR_FFT32], ids=["fft64", "fft32"])
@pytest.mark.parametrize("n", [128, 255, 384, 700, 1023, 1200, 1280, 1537, 2048, 2500, 3000, 3100, 4097, 5000, 6000, 6144])
def test_xcorr_fft_every_transform_length(eng, n, method):
    """One window length per transform size (512 ... 12288 points), odd channel and pair counts."""
    rng = np.random.default_rng(n)
    c, count = 5, 3
    hop = n // 2
    t = n + hop * (count - 1) + 5
    s = rng.standard_normal(t + 64)
    x = np.stack([s[k * 3:k * 3 + t] + 0.5 * rng.standard_normal(t) for k in range(c)])
    ref_lag, ref_cmax = _oracle_xcorr(x, n, hop, count)
    lag, cmax = eng.xcorr(x[None], [n], [hop], [count], method=method)
    assert np.array_equal(lag, ref_lag)
    assert np.allclose(cmax, ref_cmax, rtol=0, atol=1e-12, equal_nan=True)
    lag2, _ = eng.xcorr(x[None], [n], [hop], [count], max_lag=n // 10, method=method)
    ref2, _ = _oracle_xcorr(x, n, hop, count, max_lag=n // 10)
    assert np.array_equal(lag2, ref2)


@pytest.mark.parametrize("n", [1200, 3000, 4400])
def test_two_stage_redoes_single_pair_windows(eng, n):
    """AUTO's two-stage search with pair-windows on both sides of the Cauchy-Schwarz proof: most channels carry the same
    signal at small lags (stage A proves them), one channel is delayed far beyond stage A's lag bound, one is pure noise
    and one is dead, so stage B runs in item mode on a ragged list of pair-windows (odd counts, items of different windows
    sharing a packed transform).  Lags must equal the time-domain oracle's bit for bit; AUTO's cmax must agree with the
    single-stage FP32 path's to the last bits (only the window energies are summed by kernels of different widths)."""
    rng = np.random.default_rng(n + 1)
    c, count = 7, 5
    hop = n // 2
    t = n + hop * (count - 1)
    far = int(0.45 * n)                        # beyond stage A's lag bound k0 = L_A - N (336 / 1096 / 1744 here)
    s = rng.standard_normal(t + far + 64)
    x = np.stack([s[far + 2 * k:far + 2 * k + t] + 0.2 * rng.standard_normal(t) for k in range(c)])
    x[3] = s[:t] + 0.2 * rng.standard_normal(t)          # delayed by `far` samples against the others
    x[4] = rng.standard_normal(t)                        # incoherent
    x[5] = 0.0                                           # dead channel: NaN cmax, lag -(N-1)
    x[6, hop:hop + n // 3] = 0.0                          # a gap in one channel of the middle windows
    ref_lag, ref_cmax = _oracle_xcorr(x, n, hop, count)
    eng.xcorr_stats(reset=True)
    lag, cmax = eng.xcorr(x[None], [n], [hop], [count], method=_capi.XCORR_AUTO)
    st = eng.xcorr_stats(reset=True)
    assert np.array_equal(lag, ref_lag)
    assert np.allclose(cmax, ref_cmax, rtol=0, atol=1e-12, equal_nan=True)
    assert st["two_stage_windows"] == count and st["two_stage_redone"] == count
    assert np.abs(lag).max() >= far - 16                 # the far lags were found (only stage B can see them)
    lag1, cmax1 = eng.xcorr(x[None], [n], [hop], [count], method=_capi.XCORR_FFT32)
    assert np.array_equal(lag, lag1)
    assert np.allclose(cmax, cmax1, rtol=0, atol=1e-14, equal_nan=True)


def test_two_contexts_run_different_jobs_concurrently(eng):
    """Two contexts on one GPU, driven from two host threads with jobs of very different shared-memory needs (the
    two-records-in-flight mode of bench.py's e2e leg): each must return exactly what it returns alone."""
    import threading
    from narrow_band_least_squares_b200.engine import Engine
    rng = np.random.default_rng(77)
    jobs = []
    for n, c in ((3000, 6), (700, 9)):
        hop, count = n // 2, 6
        t = n + hop * (count - 1)
        s = rng.standard_normal(t + 64)
        x = np.stack([s[2 * k:2 * k + t] + 0.4 * rng.standard_normal(t) for k in range(c)])
        x[c - 1] = rng.standard_normal(t)                # an incoherent channel: stage B (item mode) has work
        jobs.append((x[None], [n], [hop], [count]))
    alone = [eng.xcorr(*j, method=_capi.XCORR_AUTO) for j in jobs]
    eng2 = Engine(0)
    try:
        engines, out, errs = (eng, eng2), [None, None], []

        def work(i):
            try:
                for _ in range(8):
                    out[i] = engines[i].xcorr(*jobs[i], method=_capi.XCORR_AUTO)
                    assert np.array_equal(out[i][0], alone[i][0])
                    assert np.array_equal(out[i][1], alone[i][1], equal_nan=True)
            except BaseException as exc:                 # noqa: BLE001 - reported by the main thread
                errs.append(exc)

        th = [threading.Thread(target=work, args=(i,)) for i in range(2)]
        for t_ in th:
            t_.start()
        for t_ in th:
            t_.join()
        assert not errs, errs
    finally:
        eng2.close()


def test_cheby1_branch_matches_reference(eng):
    """FILTER_TYPE='cheby1' (helpers.py:129-137 of the reference: one causal sosfilt pass, then the 1 % taper) through
    the package's filter_data on the GPU, against samples the unmodified reference produced (tests/golden/helpers.npz)."""
    from narrow_band_least_squares_b200 import helpers
    g = np.load(os.path.join(GOLD, "helpers.npz"), allow_pickle=True)
    seed, nch, npts, fs, ap = g["cheby_seed"]
    rec = plane_wave_record(seed=int(seed), nchans=int(nch), npts=int(npts), fs=float(fs), aperture_km=float(ap))
    for ci in range(2):
        fmin, fmax, order, ripple = g[f"cheby_{ci}_meta"]
        stf, fs_out, sos = helpers.filter_data((rec.data, rec.fs, None), "cheby1", float(fmin), float(fmax), int(order),
                                               float(ripple))
        filt = stf[0]                      # tuple in -> tuple out (stream.like_stream)
        assert np.array_equal(np.asarray(sos), g[f"cheby_{ci}_sos"])
        tol = 1e-12 * float(g[f"cheby_{ci}_rms"].max())
        assert np.allclose(filt[:, :128], g[f"cheby_{ci}_head"], rtol=0, atol=tol)
        assert np.allclose(filt[:, 4500:4628], g[f"cheby_{ci}_mid"], rtol=0, atol=tol)
        assert np.allclose(filt[:, -128:], g[f"cheby_{ci}_tail"], rtol=0, atol=tol)


def test_xcorr_fp32_screen_statistics(eng):
    """The FP32 screen hands ~1 candidate per pair-window to the exact FP64 evaluation and its maximum agrees with the
    exact one far inside the candidate band (2e-5): the margin that makes the integer lags FP64-exact."""
    rng = np.random.default_rng(11)
    n, hop, count, c = 3000, 1500, 12, 7
    t = n + hop * (count - 1)
    s = rng.standard_normal(t + 64)
    x = np.stack([s[k * 2:k * 2 + t] + 0.3 * rng.standard_normal(t) for k in range(c)])
    eng.xcorr_stats(reset=True)
    lag, cmax = eng.xcorr(x[None], [n], [hop], [count], method=_capi.XCORR_FFT32)
    st = eng.xcorr_stats(reset=True)
    ref_lag, ref_cmax = _oracle_xcorr(x, n, hop, count)
    assert np.array_equal(lag, ref_lag)
    assert np.allclose(cmax, ref_cmax, rtol=0, atol=1e-13)
    items = count * c * (c - 1) // 2
    assert items <= st["verified"] <= 1.5 * items, st
    assert st["time_domain"] == 0, st
    assert 0.0 < st["max_screen_err"] < 2e-6, st


def test_errors_mirror_reference(eng):
    from narrow_band_least_squares_b200 import ltsva
    x = np.random.default_rng(0).standard_normal((2, 4000))
    with pytest.raises(RuntimeError):
        ltsva((x, 40.0), None, None, 10.0, 0.5, rij=np.array([[0.0, 1.0], [0.0, 0.5]]), engine=eng)
    x = np.random.default_rng(0).standard_normal((3, 4000))
    with pytest.raises(RuntimeError):     # LTS needs a co-array of >= 4
        ltsva((x, 40.0), None, None, 10.0, 0.5, 0.75, rij=np.array([[0.0, 1.0, 0.3], [0.0, 0.5, 0.9]]), engine=eng)
    with pytest.raises(RuntimeError):     # collinear array: rank-deficient co-array
        ltsva((x, 40.0), None, None, 10.0, 0.5, rij=np.array([[0.0, 1.0, 2.0], [0.0, 1.0, 2.0]]), engine=eng)


@pytest.mark.parametrize("alpha", [1.0, 0.75])
def test_time_sharded_run_equals_unsharded(eng, alpha):
    """Three logical time shards (filter halos, global taper index) on one GPU reproduce the unsharded lags, LTS flags
    and estimates: the multi-GPU path without a cluster."""
    from narrow_band_least_squares_b200.engine import Plan, Results, window_grid
    from narrow_band_least_squares_b200.sharding import Shard, filter_halo, merge_band_major
    rec = plane_wave_record(seed=21, nchans=6, npts=240000, fs=100.0, aperture_km=0.5, flip_channel=3)
    fl = orc.get_freqlist_log(0.5, 5.0, 3)
    sos = np.stack([orc.design_sos("butter", fl[b], fl[b + 1], rec.fs, 2) for b in range(3)])
    wl = [40.0, 30.0, 20.0]
    grids = [window_grid(rec.npts, rec.fs, w, 0.5) for w in wl]
    n = [g[0] for g in grids]; hop = [g[1] for g in grids]; cnt = [g[2] for g in grids]
    whole = Plan(rec.nchans, rec.npts, rec.fs, sos, True, n, hop, cnt, rec.rij, alpha=alpha)
    ref = eng.run(whole, rec.data, Results(whole, pinned=False))
    halo = filter_halo(sos)
    parts, counts = {k: [] for k in ("lag", "vel", "baz", "mdccm", "sigma_tau", "weights")}, []
    world = 3
    for r in range(world):
        sh = Shard(rec.npts, n, hop, cnt, r, world, halo)
        plan = sh.plan(rec.nchans, rec.fs, sos, True, n, hop, rec.rij, alpha=alpha)
        res = eng.run(plan, rec.data[:, sh.seg_lo:sh.seg_hi], Results(plan, pinned=False))
        counts.append(plan.win_count)
        for k in parts:
            a = getattr(res, k)
            if a is not None:
                parts[k].append(a[:plan.nwin])
    assert np.array_equal(merge_band_major(parts["lag"], counts), ref.lag), "sharded lags differ"
    if alpha != 1.0:
        assert np.array_equal(merge_band_major(parts["weights"], counts), ref.weights), "sharded LTS flags differ"
    for k in ("vel", "sigma_tau", "mdccm"):
        assert np.allclose(merge_band_major(parts[k], counts), getattr(ref, k), rtol=1e-9, atol=1e-12, equal_nan=True), k
    d = np.abs(((merge_band_major(parts["baz"], counts) - ref.baz + 180.0) % 360.0) - 180.0)
    assert np.nanmax(d, initial=0.0) < 1e-6


def test_c3_sized_methods_agree_end_to_end(eng):
    """BASELINE config 3 at its real array and band sizes (25 elements = 300 pairs, 20 log bands, windows 60 -> 30 s, i.e.
    transforms of 12288, 10240, 8192 and 6144 points) on a 15-minute record: the default path (FP32 screen + exact
    FP64 candidates), the FP64 FFT path and the exact time-domain kernel give the same integer lags, cmax to 1e-12 and
    the same estimates; the screen's error stays far inside its candidate band and nothing needs the fallback."""
    from narrow_band_least_squares_b200.configs import config_plan
    from narrow_band_least_squares_b200.engine import Results
    from narrow_band_least_squares_b200.synth import config_record
    rec = config_record("c3", hours=0.25)
    out = {}
    for name, method in (("fft32", _capi.XCORR_FFT32), ("auto", _capi.XCORR_AUTO), ("fft64", _capi.XCORR_FFT),
                         ("direct", _capi.XCORR_DIRECT)):
        plan = config_plan("c3", rec.npts, rec.rij, method=method)
        if name in ("fft32", "auto"):
            eng.xcorr_stats(reset=True)
        res = eng.run(plan, rec.data, Results(plan, pinned=False))
        if name == "fft32":
            st = eng.xcorr_stats(reset=True)
        if name == "auto":
            st2 = eng.xcorr_stats(reset=True)
        out[name] = {k: np.array(getattr(res, k)[:plan.nwin]) for k in ("lag", "cmax", "vel", "baz", "mdccm", "sigma_tau")}
        nwin, npairs = plan.nwin, plan.npairs
    assert npairs == 300 and nwin > 600
    # two-stage search of AUTO: (nearly) every window went through stage A - a band where most windows fail hands the next
    # three bands to the single-stage search - and the noise-only minute (and little else) needed stage B
    assert 0.5 * nwin < st2["two_stage_windows"] <= nwin and 0 < st2["two_stage_redone"] < 0.5 * nwin, st2
    for other in ("auto", "fft64", "direct"):
        assert np.array_equal(out["fft32"]["lag"], out[other]["lag"]), other
        assert np.allclose(out["fft32"]["cmax"], out[other]["cmax"], rtol=0, atol=1e-12, equal_nan=True), other
        for k in ("vel", "baz", "mdccm", "sigma_tau"):
            assert np.allclose(out["fft32"][k], out[other][k], rtol=1e-9, atol=1e-12, equal_nan=True), (other, k)
    assert st["time_domain"] == 0 and st["max_screen_err"] < 5e-6, st
    assert nwin * npairs <= st["verified"] <= 1.2 * nwin * npairs, st
    v = out["fft32"]["vel"]
    good = np.isfinite(v)
    assert good.mean() > 0.5 and abs(np.median(v[good]) - 0.34) < 0.02      # the synthetic plane wave is recovered
