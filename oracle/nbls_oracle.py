"""CPU oracle for the NBLS hot path — a numpy/scipy restatement of the reference algorithm.

TEST INFRASTRUCTURE ONLY.  Nothing in the product package
(``narrow_band_least_squares_b200/``) imports this module; only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / ``--impl reference``
legs may.  It is the checker, never the thing shipped or measured as the product.

Pinning: the reference has no tests or golden vectors of its
own, so this restatement is pinned against outputs of the UNMODIFIED reference
executed in the build container (``oracle/run_reference.py`` ->
``tests/golden/*.npz``; checked by ``tests/test_oracle_golden.py``).
Exception — **parity unpinned** for the band-pass/taper arithmetic and the
Vincenty inverse: those live in obspy, which the reference neither vendors nor
pins (environment.yml:9); both this file and the import stub restate obspy's
documented behaviour with scipy.

All citations are relative to /root/reference/.
"""
from __future__ import annotations

import math
import os

import numpy as np
from scipy import signal
from scipy.linalg import lstsq as sp_lstsq
from scipy.special import erfinv

try:  # the reference JIT-compiles its LTS loop with numba (lts_classes.py:16,39,58)
    from numba import njit
    HAVE_NUMBA = os.environ.get("NBLS_ORACLE_NO_NUMBA", "0") != "1"
except Exception:  # pragma: no cover
    HAVE_NUMBA = False

if not HAVE_NUMBA:  # pragma: no cover
    def njit(*a, **k):
        if len(a) == 1 and callable(a[0]):
            return a[0]
        return lambda f: f


# --------------------------------------------------------------------------------------
# K1: band-pass + taper            helpers.py:108-141  (+ obspy, restated: SURVEY A.1/A.3)
# --------------------------------------------------------------------------------------
def design_sos(ftype, fmin, fmax, fs, order, ripple=None):
    """Second-order sections for one band.

    'butter' follows obspy.signal.filter.bandpass as called at helpers.py:127
    (normalised corner frequencies -> zpk -> zpk2sos); 'cheby1' follows
    helpers.py:130 (scipy iirfilter with fs=, output='sos').
    """
    if ftype == "butter":
        fe = 0.5 * fs
        lo, hi = fmin / fe, fmax / fe
        if hi - 1.0 > -1e-6:
            z, p, k = signal.iirfilter(order, lo, btype="highpass", ftype="butter", output="zpk")
        else:
            if lo > 1:
                raise ValueError("Selected low corner frequency is above Nyquist.")
            z, p, k = signal.iirfilter(order, [lo, hi], btype="band", ftype="butter", output="zpk")
        return signal.zpk2sos(z, p, k)
    if ftype == "cheby1":
        return signal.iirfilter(order, [fmin, fmax], rp=ripple, btype="band", analog=False,
                                ftype="cheby1", fs=fs, output="sos")
    raise ValueError(ftype)


def taper_weights(npts, max_percentage=0.01):
    """Hann ramp halves of obspy's Trace.taper(max_percentage), helpers.py:139."""
    wlen = min(int(max_percentage * npts), int(npts / 2))
    w = signal.windows.hann(2 * wlen if 2 * wlen == npts else 2 * wlen + 1)
    return wlen, w[:wlen].copy(), w[len(w) - wlen:].copy()


def filter_record(data, fs, ftype, fmin, fmax, order, ripple=None, taper=0.01):
    """(C, T) float64 -> filtered+tapered (C, T).  helpers.py:108-141."""
    data = np.asarray(data, dtype=np.float64)
    sos = design_sos(ftype, fmin, fmax, fs, order, ripple)
    out = np.empty_like(data)
    for c in range(data.shape[0]):
        if ftype == "butter":          # zerophase=True: forward, then on the reversed signal; zero state, no padding
            y = signal.sosfilt(sos, data[c])
            y = signal.sosfilt(sos, y[::-1])[::-1]
        else:                          # cheby1: one causal pass (helpers.py:131-137)
            y = signal.sosfilt(sos, data[c])
        out[c] = y
    if taper:
        wlen, up, down = taper_weights(data.shape[1], taper)
        if wlen:
            out[:, :wlen] *= up
            out[:, data.shape[1] - wlen:] *= down
    return out, sos


# --------------------------------------------------------------------------------------
# window grid and co-array      lts_data_class.py:58-66, lts_classes.py:836-850
# --------------------------------------------------------------------------------------
def window_grid(npts, fs, window_length, window_overlap):
    n = int(window_length * fs)
    hop = int(np.round((1 - window_overlap) * n))
    starts = np.arange(0, npts - n, hop, dtype="int")
    return n, hop, starts


def co_array(rij):
    c = rij.shape[1]
    pairs = [(i, j) for i in range(c - 1) for j in range(i + 1, c)]
    ia = np.array([p[0] for p in pairs])
    ib = np.array([p[1] for p in pairs])
    xij = (rij[:, ia] - rij[:, ib]).T
    mad = np.array([1.4826 * np.median(np.abs(xij[:, d])) for d in range(2)])
    return pairs, xij, mad, xij / mad


# --------------------------------------------------------------------------------------
# K2: windowed pairwise cross-correlation       lts_classes.py:852-880
# --------------------------------------------------------------------------------------
def correlate_windows(data_tc, starts, n, fs, pairs, tvec=None, max_lag=None):
    """data_tc: (T, C).  Returns lag (P,W) int32 [lag = argmax - (N-1)], cmax (P,W), tau (P,W),
    mdccm (W,), mad (W,), t (W,).

    ``max_lag`` (not in the reference) restricts the argmax to |lag| <= max_lag by slicing
    rows N-1-L .. N-1+L of cij — used only to check the product's restricted-lag mode.
    """
    p, w = len(pairs), len(starts)
    lag = np.zeros((p, w), dtype=np.int32)
    cmax = np.full((p, w), np.nan)
    tau = np.zeros((p, w))
    mdccm = np.full(w, np.nan)
    mad = np.zeros(w)
    t = np.full(w, np.nan)
    with np.errstate(invalid="ignore", divide="ignore"):
        for jj, t0 in enumerate(starts):
            seg = data_tc[t0:t0 + n]
            if tvec is not None:
                t[jj] = tvec[t0 + int(n / 2)]
            energy = np.sum(seg * seg, axis=0)
            for k, (a, b) in enumerate(pairs):
                xa, xb = np.ascontiguousarray(seg[:, a]), np.ascontiguousarray(seg[:, b])
                c = np.correlate(xa, xb, mode="full") / np.sqrt(np.sum(xa * xa) * np.sum(xb * xb))
                if max_lag is not None and max_lag < n - 1:
                    lo = n - 1 - max_lag
                    m = int(np.argmax(c[lo:n + max_lag])) + lo
                else:
                    m = int(np.argmax(c))
                lag[k, jj] = m - (n - 1)
                cmax[k, jj] = c[m]
            del energy
            col = cmax[:, jj]
            mdccm[jj] = np.nanmedian(col) if not np.all(np.isnan(col)) else np.nan
            # delay = argmax + 1 ; tau = (N - delay)/fs     (lts_classes.py:875-876)
            tau[:, jj] = (n - (lag[:, jj] + n)) / fs
            mad[jj] = 1.4826 * np.median(np.abs(tau[:, jj]))
    return lag, cmax, tau, mdccm, mad, t


# --------------------------------------------------------------------------------------
# K3a: OLS                           lts_classes.py:886-936
# --------------------------------------------------------------------------------------
def slowness_to_vel_baz(zx, zy):
    vel = 1.0 / math.hypot(zx, zy) if (zx != 0.0 or zy != 0.0) else np.inf
    baz = (math.atan2(zx, zy) * 180.0 / math.pi - 360.0) % 360.0
    return vel, baz


def ols_solve(tau, mad, xij, xij_mad, xij_std):
    p, w = tau.shape
    vel = np.full(w, np.nan)
    baz = np.full(w, np.nan)
    sig = np.full(w, np.nan)
    z_all = np.full((2, w), np.nan)
    q, r = np.linalg.qr(xij_std)
    for jj in range(w):
        if mad[jj] == 0 or np.count_nonzero(tau[:, jj]) < p - 2:
            continue
        y = tau[:, jj] / mad[jj]
        z = sp_lstsq(r, q.conj().T @ y)[0]
        z = z * (mad[jj] / xij_mad)
        z_all[:, jj] = z
        vel[jj] = 1.0 / np.linalg.norm(z, 2)
        baz[jj] = (np.arctan2(z[0], z[1]) * 180 / np.pi - 360) % 360
        resid = tau[:, jj] - xij @ z
        with np.errstate(invalid="ignore"):
            sig[jj] = np.sqrt(tau[:, jj] @ resid / (p - 2))
    return vel, baz, sig, z_all


# --------------------------------------------------------------------------------------
# K3b: FAST-LTS                      lts_classes.py:16-150, 156-344, 643-793, 971-1036
# --------------------------------------------------------------------------------------
def lcg_subsets(p, count, npar=2):
    """The reference's 16-bit LCG subset stream (lts_classes.py:16-36), restarted at seed 0.

    Returns (count, npar) int64 indices and the seed after each draw.  Integer arithmetic is
    exact: seed*5761+999 < 2**53, so the float floor/divide of the reference equals this.
    """
    out = np.zeros((count, npar), dtype=np.int64)
    seeds = np.zeros(count, dtype=np.int64)
    seed = 0
    for i in range(count):
        got = []
        for _ in range(npar):
            while True:
                seed = (seed * 5761 + 999) % 65536
                num = int(math.floor((seed / 65536) * p))
                if num not in got:
                    break
            got.append(num)
        out[i] = got
        seeds[i] = seed
    return out, seeds


def h_value(p, alpha, ndim=2):
    """lts_classes.py:1013."""
    half = np.floor((p + ndim + 1) / 2)
    return int(np.floor(2 * half - p + 2 * (p - half) * alpha))


def _small_sample_factor(coef500, coef875, p, n, alpha):
    """Pison et al. (2002) finite-sample interpolation shared by lts_classes.py:156-218 and :260-323."""
    def fit(co):
        # keep the reference's operation order: y = 1 + a/p^b, then log(1 - y)
        y1 = 1 + co[0][0] / np.power(p, co[1][0])
        y2 = 1 + co[0][1] / np.power(p, co[1][1])
        y = np.array([np.log(1 - y1), np.log(1 - y2)])
        xm = np.array([[1.0, np.log(1.0 / (co[2][0] * p ** 2))],
                       [1.0, np.log(1.0 / (co[2][1] * p ** 2))]])
        c = np.linalg.lstsq(xm, y.reshape(2, 1), rcond=-1)[0]
        return 1.0 - np.exp(c[0]) / np.power(n, c[1])
    f500, f875 = fit(coef500), fit(coef875)
    if 0.5 <= alpha <= 0.875:
        f = f500 + (f875 - f500) / 0.375 * (alpha - 0.5)
    elif 0.875 < alpha < 1:
        f = f875 + (1 - f875) / 0.125 * (alpha - 0.875)
    else:
        raise ValueError("alpha must lie in [0.5, 1)")
    return float(np.ndarray.item(1.0 / f))


def raw_corfactor(p, n, alpha):
    return _small_sample_factor(
        [[-0.487338281979106, -0.340762058011], [0.405511279418594, 0.37972360544988], [3, 5]],
        [[-0.251778730491252, -0.146660023184295], [0.883966931611758, 0.86292940340761], [3, 5]],
        p, n, alpha)


def rew_corfactor(p, n, alpha):
    return _small_sample_factor(
        [[-0.417574780492848, -0.175753709374146], [1.83958876341367, 1.8313809497999], [3, 5]],
        [[-0.267522855927958, -0.161200683014406], [1.17559984533974, 1.21675019853961], [3, 5]],
        p, n, alpha)


def raw_consfactor(h, n):
    """lts_classes.py:221-245."""
    x = (h + n) / (2 * n)
    q = np.sqrt(2) * erfinv(2 * x - 1)
    phi = (1 / np.sqrt(2 * np.pi)) * np.exp(-0.5 * q ** 2)
    return 1 / np.sqrt(1 - (2 * n / (h * (1 / q))) * phi)


def rew_consfactor(nw, n):
    """lts_classes.py:326-344 as a function of the number of unit weights ``nw``."""
    with np.errstate(all="ignore"):
        q = erfinv(2 * ((nw + n) / (2 * n)) - 1) * np.sqrt(2)
        a = (1 / np.sqrt(2 * np.pi)) * np.exp(-0.5 * (1 / (1 / q)) ** 2)
        b = 1 / q
        return 1 / np.sqrt(1 - ((2 * n) / (nw * b)) * a)


@njit(cache=False)
def _ls2(xs, ys):
    """Least squares through QR, the reference's three-call idiom (lts_classes.py:84-86)."""
    q, r = np.linalg.qr(xs)
    return np.linalg.lstsq(r, q.conj().T @ ys)[0]


@njit(cache=False)
def _csteps(xv, yv, z, h, limit):
    """Concentration steps until the objective repeats (lts_classes.py:89-103 / :125-139)."""
    resid = yv - xv @ z
    prev = 0.0
    obj = 0.0
    for kk in range(limit):
        order = np.argsort(np.abs(resid).flatten())
        keep = order.flatten()[0:h]
        z = _ls2(xv[keep, :], yv[keep])
        resid = yv - xv @ z
        srt = np.sort(np.abs(resid).flatten())
        obj = np.sum(srt[0:h] ** 2)
        if kk >= 1 and obj == prev:
            break
        prev = obj
    return z, obj


@njit(cache=False)
def _fast_lts_window(yv, xv, subsets, h, n_starts, n_keep, steps1, steps2):
    """One window of FAST-LTS on standardised data (lts_classes.py:70-142). Returns (2,1) coefficients."""
    best_obj = np.full(n_keep, np.inf)
    best_z = np.full((2, n_keep), np.nan)
    used = 0
    for ii in range(n_starts):
        idx = subsets[used]
        used += 1
        z = _ls2(xv[idx, :], yv[idx])
        z, obj = _csteps(xv, yv, z, h, steps1)
        if obj < np.max(best_obj):
            # keep the n_keep smallest, duplicates included, arrival order on ties
            # (check_array lts_classes.py:39-55: its duplicate test never fires, SURVEY a-7)
            allobj = np.concatenate((best_obj, np.array([obj])))
            allz = np.concatenate((best_z, z), axis=1)
            sel = np.argsort(allobj)[0:n_keep]
            best_obj = allobj[sel]
            best_z = allz[:, sel]
    final_obj = np.inf
    final_z = np.full((2, 1), np.nan)
    for ii in range(n_keep):
        if np.isfinite(best_obj[ii]):
            z = best_z[:, ii].copy().reshape((2, 1))
        else:
            idx = subsets[used]
            used += 1
            z = _ls2(xv[idx, :], yv[idx])
        if np.isfinite(z[0, 0]):
            z, obj = _csteps(xv, yv, z, h, steps2)
            if obj < final_obj:
                final_obj = obj
                final_z = z.copy()
    return final_z


def fast_lts(tau, mad, xij_std, xij_mad, h, n_starts=500, n_keep=10, steps1=4, steps2=100):
    """lts_classes.py:58-150. tau (P,W). Returns raw slowness coefficients (2,W), NaN where skipped."""
    p, w = tau.shape
    subsets, _ = lcg_subsets(p, n_starts + n_keep)
    coeffs = np.full((2, w), np.nan)
    xv = np.ascontiguousarray(xij_std)
    for jj in range(w):
        if mad[jj] == 0 or np.count_nonzero(tau[:, jj]) < p - 2:
            continue
        yv = np.ascontiguousarray((tau[:, jj] / mad[jj]).reshape(p, 1))
        z = _fast_lts_window(yv, xv, subsets, h, n_starts, n_keep, steps1, steps2)
        coeffs[:, jj] = z[:, 0] * mad[jj] / xij_mad
    return coeffs


def lts_post(tau, xij, coeffs, alpha, h):
    """lts_classes.py:643-764 without the uncertainty ellipse (:715-720,:732-759)."""
    p, w = tau.shape
    raw = raw_corfactor(2, p, alpha) * raw_consfactor(h, p)
    rew1 = rew_corfactor(2, p, alpha)
    quant = 2.2414027276049473
    vel = np.full(w, np.nan)
    baz = np.full(w, np.nan)
    sig = np.full(w, np.nan)
    weights = np.zeros((p, w), dtype=np.uint8)
    zf = np.full((2, w), np.nan)
    for jj in range(w):
        if np.count_nonzero(tau[:, jj]) < p - 2:
            continue
        y = tau[:, jj].reshape(p, 1)
        resid = y - xij @ coeffs[:, jj].reshape(2, 1)
        s0 = np.sqrt(np.sum(np.sort(resid.flatten() ** 2)[0:h]) / h) * raw
        if np.abs(s0) < 1e-7:
            # the reference's branch here raises IndexError at lts_classes.py:717;
            # the oracle defines: exact fit, keep the raw coefficients.
            wt = (np.abs(resid) < 1e-7).flatten()
            z = coeffs[:, jj].reshape(2, 1)
        else:
            wt = (np.abs(resid / s0) <= quant).flatten()
            q, r = np.linalg.qr(xij[wt, :])
            z = np.linalg.lstsq(r, q.conj().T @ y[wt], rcond=None)[0]
            resid = y - xij @ z
            nw = int(np.sum(wt))
            with np.errstate(all="ignore"):
                scale = np.sqrt(np.sum(resid[wt] ** 2) / (nw - 1)) * rew1
                if nw != p:
                    scale *= rew_consfactor(nw, p)
                wt = (np.abs(resid / scale) <= 2.5).flatten()
        zf[:, jj] = z[:, 0]
        vel[jj] = 1 / np.linalg.norm(z, 2)
        baz[jj] = (np.arctan2(z[0][0], z[1][0]) * 180 / np.pi - 360) % 360
        r2 = tau[wt, jj] - xij[wt, :] @ z[:, 0]
        m_w = int(np.sum(wt))
        with np.errstate(all="ignore"):
            val = np.float64(tau[wt, jj] @ r2) / np.float64(m_w - 2)
            if not (np.isnan(val) or val < 0):      # 'invalid' raises in the reference and is swallowed (:725-730)
                sig[jj] = np.sqrt(val)
        weights[:, jj] = wt
    return vel, baz, sig, weights, zf


def flagged_elements(weight_col, pairs):
    """lts_classes.py:767-793: 1-based element numbers of zero-weight pairs, firsts then seconds."""
    bad = np.where(weight_col == 0)[0]
    a = np.array([p[0] for p in pairs]) + 1
    b = np.array([p[1] for p in pairs]) + 1
    return np.concatenate((a[bad], b[bad]))


def build_stdict(weights, t, pairs, nchans):
    """lts_classes.py:1028-1036."""
    d = {}
    for jj in range(weights.shape[1]):
        s = flagged_elements(weights[:, jj], pairs)
        if len(s) > 0:
            d[str(t[jj])] = s
        if jj == weights.shape[1] - 1:
            d["size"] = nchans
    return d


# --------------------------------------------------------------------------------------
# ltsva and the NBLS drivers        ltsva.py:9-54, narrow_band_least_squares.py:8-323
# --------------------------------------------------------------------------------------
def ltsva_oracle(data_ct, fs, tvec, rij, window_length, window_overlap, alpha=1.0, max_lag=None):
    """data_ct: filtered (C,T).  Returns a dict of every stage's outputs."""
    data_ct = np.asarray(data_ct, dtype=np.float64)
    c, npts = data_ct.shape
    if c < 3:
        raise RuntimeError("The array must have at least 3 elements for well-posed least squares estimation.")
    n, hop, starts = window_grid(npts, fs, window_length, window_overlap)
    pairs, xij, xij_mad, xij_std = co_array(rij)
    if np.linalg.matrix_rank(xij) < 2:
        raise RuntimeError("Co-array is ill posed for the least squares problem.")
    data_tc = np.ascontiguousarray(data_ct.T)
    lag, cmax, tau, mdccm, mad, t = correlate_windows(data_tc, starts, n, fs, pairs, tvec, max_lag)
    out = dict(n=n, hop=hop, starts=starts, pairs=pairs, xij=xij, lag=lag, cmax=cmax, tau=tau, mdccm=mdccm,
               mad=mad, t=t)
    if alpha == 1.0:
        vel, baz, sig, z = ols_solve(tau, mad, xij, xij_mad, xij_std)
        out.update(vel=vel, baz=baz, sigma_tau=sig, z=z, stdict={})
    else:
        if len(pairs) < 4:
            raise RuntimeError("The co-array must have at least 4 elements for least trimmed squares.")
        h = h_value(len(pairs), alpha)
        coeffs = fast_lts(tau, mad, xij_std, xij_mad, h)
        vel, baz, sig, wts, zf = lts_post(tau, xij, coeffs, alpha, h)
        out.update(vel=vel, baz=baz, sigma_tau=sig, z=zf, slowness_coeffs=coeffs, element_weights=wts, h=h,
                   stdict=build_stdict(wts, t, pairs, c))
    return out


def get_freqlist_log(fmin, fmax, nbands):
    """helpers.py:29-34."""
    return np.logspace(math.log(fmin, 10), math.log(fmax, 10), num=nbands + 1)


def nbls_vector_len(npts, fs, winlen_list, winover):
    """narrow_band_least_squares.py:41-47 (the reference's sizing quirk, SURVEY a-14)."""
    sampinc = int((1 - winover) * winlen_list[-1])
    return int((len(np.arange(0, npts, sampinc)) - 1) / fs)


def nbls_band(ii, data_ct, fs, tvec, rij, freqlist, winlen_list, winover, alpha, ftype, order, ripple,
              max_lag=None):
    """One band of the driver loop (narrow_band_least_squares.py:67-124 / narrow_band_loop :134-218)."""
    filt, sos = filter_record(data_ct, fs, ftype, freqlist[ii], freqlist[ii + 1], order, ripple)
    res = ltsva_oracle(filt, fs, tvec, rij, winlen_list[ii], winover, alpha, max_lag)
    res["sos"] = sos
    return res


def nbls_oracle(data_ct, fs, tvec, rij, freqlist, winlen_list, winover, alpha, ftype="butter", order=2,
                ripple=0.01, n_jobs=1, max_lag=None):
    """The driver: per-band filter + ltsva, padded into (B, vector_len) arrays
    (narrow_band_least_squares.py:8-127; n_jobs != 1 mirrors the joblib twin :223-323).
    Rows beyond num_compute are NaN here (uninitialised in the reference)."""
    nb = len(freqlist) - 1
    npts = data_ct.shape[1]
    vlen = nbls_vector_len(npts, fs, winlen_list, winover)
    if n_jobs == 1:
        bands = [nbls_band(ii, data_ct, fs, tvec, rij, freqlist, winlen_list, winover, alpha, ftype, order,
                           ripple, max_lag) for ii in range(nb)]
    else:
        from joblib import Parallel, delayed
        bands = Parallel(n_jobs=n_jobs)(delayed(nbls_band)(ii, data_ct, fs, tvec, rij, freqlist, winlen_list,
                                                           winover, alpha, ftype, order, ripple, max_lag)
                                        for ii in range(nb))
    arr = {k: np.full((nb, vlen), np.nan) for k in ("vel", "baz", "mdccm", "t", "sigma_tau")}
    ncomp = []
    stdict_all = None if alpha == 1.0 else {}
    for ii, r in enumerate(bands):
        w = len(r["vel"])
        ncomp.append(w)
        for k in arr:
            arr[k][ii, :w] = r[k]
        if alpha != 1.0:
            for key, val in r["stdict"].items():
                stdict_all[key if key == "size" else f"{ii + 1:02d}_{key}"] = val
    return dict(arr, num_compute=ncomp, stdict_all=stdict_all, bands=bands)
