#!/usr/bin/env python
"""Counts integer-lag mismatches between the GPU path and the reference algorithm on a LONG slice of the benchmark
workload (default: 1 h of c3 = 25 elements, 20 bands, ~3 400 windows, ~1.0 M pair-windows) — the count the bench line's
3-minute `parity` block is too short to give.

The reference's correlate (lts_classes.py:867-875: np.correlate 'full' + argmax) is O(N^2) per pair-window, hours of CPU
for this many; so the CPU side screens with a float64 FFT correlation of the oracle-filtered samples (same lags except
on ~1e-13 near-ties) and every pair-window where GPU lag != FFT arg-max is ARBITRATED by np.correlate itself — on the
oracle-filtered samples (what the reference would return) and on the GPU-filtered samples (which tells a filter-rounding
near-tie from a cross-correlation error).  Prints one JSON object (kept as profiles/r2_parity_count.json).

  python scripts/parity_count.py [hours=1.0] [workload=c3]        (GPU box; uses all host cores)
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from narrow_band_least_squares_b200.configs import config_job, config_plan  # noqa: E402
from narrow_band_least_squares_b200.engine import Engine, Results  # noqa: E402
from narrow_band_least_squares_b200.synth import config_record  # noqa: E402
from oracle import nbls_oracle as orc  # noqa: E402


def band_lags_fft(filt, n, hop, nwin, pairs):
    """argmax lags of every (window, pair) of one band by float64 FFT correlation; returns (W, P) int32."""
    L = 1
    while L < 2 * n:
        L *= 2
    ia = np.array([p[0] for p in pairs]); ib = np.array([p[1] for p in pairs])
    out = np.empty((nwin, len(pairs)), dtype=np.int32)
    for w in range(nwin):
        seg = filt[:, w * hop:w * hop + n]
        X = np.fft.rfft(seg, L, axis=1)
        c = np.fft.irfft(X[ia] * np.conj(X[ib]), L, axis=1)            # c[k] = sum_n a[n+k] b[n], circular
        full = np.concatenate((c[:, L - (n - 1):], c[:, :n]), axis=1)    # lags -(n-1) .. n-1
        en = np.sum(seg * seg, axis=1)
        dead = (en[ia] * en[ib]) == 0
        lag = np.argmax(full, axis=1) - (n - 1)
        lag[dead] = -(n - 1)
        out[w] = lag
    return out


def exact_lag(xa, xb):
    n = len(xa)
    with np.errstate(invalid="ignore", divide="ignore"):
        c = np.correlate(xa, xb, mode="full") / np.sqrt(np.sum(xa * xa) * np.sum(xb * xb))
    return int(np.argmax(c)) - (n - 1)


def one_band(b, data, fs, j, gpu_lag):
    filt, _ = orc.filter_record(data, fs, j["filter_type"], j["freqlist"][b], j["freqlist"][b + 1], j["filter_order"],
                                j["filter_ripple"])
    n, hop, starts = orc.window_grid(data.shape[1], fs, j["winlen_list"][b], j["winover"])
    pairs = [(i, k) for i in range(data.shape[0] - 1) for k in range(i + 1, data.shape[0])]
    nwin = len(starts)
    cpu = band_lags_fft(filt, n, hop, nwin, pairs)
    bad = np.argwhere(cpu != gpu_lag)
    res = dict(band=b, n=n, hop=hop, windows=nwin, pair_windows=int(nwin * len(pairs)),
               fft_screen_disagreements=int(len(bad)), mismatches_vs_np_correlate=0, of_which_match_on_gpu_filtered_samples=0,
               open_items=[])
    for w, p in bad:
        a, c = pairs[p]
        t0 = w * hop
        ref = exact_lag(np.ascontiguousarray(filt[a, t0:t0 + n]), np.ascontiguousarray(filt[c, t0:t0 + n]))
        if ref != gpu_lag[w, p]:
            res["mismatches_vs_np_correlate"] += 1
            res["open_items"].append((int(w), int(a), int(c), int(gpu_lag[w, p])))     # classified by the caller
    return res


def main():
    from joblib import Parallel, delayed
    hours = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
    name = sys.argv[2] if len(sys.argv) > 2 else "c3"
    j = config_job(name)
    rec = config_record(name, hours=hours)
    eng = Engine(0)
    plan = config_plan(name, rec.npts, rec.rij)
    t0 = time.time()
    res = eng.run(plan, rec.data, Results(plan, pinned=False, want_cmax=False))
    t_gpu = time.time() - t0
    t0 = time.time()
    rows = Parallel(n_jobs=min(os.cpu_count() or 1, j["nbands"]))(
        delayed(one_band)(b, rec.data, rec.fs, j, np.array(res.lag[plan.band_rows(b)])) for b in range(j["nbands"]))
    for r in rows:        # a true mismatch: does np.correlate on the GPU-FILTERED samples give the GPU's lag (filter rounding)?
        if r["open_items"]:
            b = r["band"]
            sos = [orc.design_sos(j["filter_type"], j["freqlist"][b], j["freqlist"][b + 1], rec.fs, j["filter_order"])]
            gf = eng.filter_bands(rec.data, rec.fs, sos)[0]
            for w, a, c, glag in r["open_items"]:
                t1 = w * r["hop"]
                alt = exact_lag(np.ascontiguousarray(gf[a, t1:t1 + r["n"]]), np.ascontiguousarray(gf[c, t1:t1 + r["n"]]))
                r["of_which_match_on_gpu_filtered_samples"] += int(alt == glag)
        r.pop("open_items")
    out = {"workload": f"{name}, {hours:g} h, {rec.nchans} elements, {j['nbands']} bands", "estimates": int(plan.nwin),
           "pair_windows": int(sum(r["pair_windows"] for r in rows)),
           "fft_screen_disagreements": int(sum(r["fft_screen_disagreements"] for r in rows)),
           "lag_mismatches_vs_np_correlate": int(sum(r["mismatches_vs_np_correlate"] for r in rows)),
           "of_which_match_on_gpu_filtered_samples": int(sum(r["of_which_match_on_gpu_filtered_samples"] for r in rows)),
           "gpu_seconds_wall": t_gpu, "cpu_seconds_wall": time.time() - t0, "per_band": rows}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
