#!/usr/bin/env python
"""Run the UNMODIFIED reference (/root/reference) on seeded synthetic records and
write golden fixtures to tests/golden/.   TEST INFRASTRUCTURE ONLY.

Runs in the build container only (the GPU box has no /root/reference).  The
reference is imported from where it lies; nothing is copied.  Two shims make
it importable here:
  * oracle/refstubs/{obspy,matplotlib}: stand-in packages (see their README);
  * numpy>=2 removed the 'complex_' alias used at
    narrow_band_least_squares.py:58-59,276-277 -> the module-local ``np`` is
    replaced by a proxy that maps that dtype string to complex128.

Usage:  python oracle/run_reference.py [--only NAME]
"""
from __future__ import annotations

import argparse
import hashlib
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.environ.get("NBLS_REFERENCE", "/root/reference")
GOLD = os.path.join(ROOT, "tests", "golden")


def import_reference():
    """The reference modules, imported from where they lie (oracle/refrun.py holds the shims)."""
    sys.path.insert(0, ROOT)
    from oracle import refrun
    if not os.path.isdir(REF):
        raise SystemExit(f"reference tree not found at {REF}")
    return refrun.import_reference(REF)


def make_stream(rec):
    from oracle import refrun
    return refrun.make_stream(rec)


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def stdict_to_arrays(stdict):
    """dict[str -> int array] -> (keys, concatenated values, offsets); 'size' kept separately."""
    keys = [k for k in stdict if k != "size"]
    vals = [np.asarray(stdict[k], dtype=np.int64) for k in keys]
    off = np.cumsum([0] + [len(v) for v in vals]).astype(np.int64)
    cat = np.concatenate(vals) if vals else np.zeros(0, np.int64)
    return np.array(keys, dtype=object).astype(str) if keys else np.zeros(0, dtype="<U1"), cat, off, int(stdict.get("size", -1))


def run_estimator(ref_cls, DataBin, st, lat, lon, winlen, winover, alpha):
    """Stage-level run (the estimator protocol of lts_classes.py:799-1036)."""
    data = DataBin(winlen, winover, alpha)
    data.build_data_arrays(st, list(lat), list(lon), None)
    est = ref_cls.OLSEstimator(data) if alpha == 1.0 else ref_cls.LTSEstimator(data)
    est.correlate(data)
    # cmax per pair-window is not kept by the reference; recompute the same expression
    # (lts_classes.py:867-873) for the fixture.
    nits, P, N = data.nits, est.co_array_num, data.winlensamp
    cmax = np.full((P, nits), np.nan)
    with np.errstate(invalid="ignore", divide="ignore"):
        for jj in range(nits):
            t0 = data.intervals[jj]
            for k, (a, b) in enumerate(est.idx_pair):
                xa = data.data[t0:t0 + N, a]
                xb = data.data[t0:t0 + N, b]
                c = np.correlate(xa, xb, mode="full") / np.sqrt(np.sum(xa * xa) * np.sum(xb * xb))
                cmax[k, jj] = c.max()
    tau = est.tau[:, :, 0].copy()
    est.solve(data)
    out = dict(rij=data.rij, intervals=data.intervals, winlensamp=data.winlensamp, sampinc=data.sampinc,
               tau=tau, lag=np.rint(-tau * data.sampling_rate).astype(np.int32), cmax=cmax,
               mdccm=est.mdccm, t=est.t, mad=est.time_delay_mad, vel=est.lts_vel, baz=est.lts_baz,
               sigma_tau=est.sigma_tau, conf_int_vel=est.conf_int_vel, conf_int_baz=est.conf_int_baz,
               xij=est.xij, xij_mad=est.xij_mad)
    if alpha != 1.0:
        k, v, o, size = stdict_to_arrays(est.stdict)
        skipped = (est.time_delay_mad == 0) | (np.count_nonzero(tau, axis=0) < P - 2)
        coeffs = est.slowness_coeffs.copy()
        coeffs[:, skipped] = np.nan          # np.empty garbage on skipped windows (lts_classes.py:978)
        out.update(element_weights=est.element_weights.astype(np.uint8), slowness_coeffs=coeffs, h=est.h,
                   stdict_keys=k, stdict_vals=v, stdict_off=o, stdict_size=size)
    return out


CASES = {
    # name: (synth kwargs, processing kwargs)
    "g1_ols_c4": (dict(seed=1001, nchans=4, npts=24000, fs=40.0, aperture_km=0.2),
                  dict(nbands=8, fmin=0.5, fmax=5.0, alpha=1.0, winlen=("constant", 30.0), winover=0.5, order=2)),
    "g2_lts_c6": (dict(seed=1002, nchans=6, npts=30000, fs=100.0, aperture_km=0.5, flip_channel=3, shift_channel=1),
                  dict(nbands=3, fmin=0.5, fmax=5.0, alpha=0.75, winlen=("constant", 30.0), winover=0.5, order=2)),
    "g3_ols_c8_adaptive_dead": (dict(seed=1003, nchans=8, npts=36000, fs=100.0, aperture_km=1.0, dead_channel=5),
                                dict(nbands=4, fmin=0.5, fmax=5.0, alpha=1.0, winlen=("adaptive", 40.0, 20.0), winover=0.5, order=2)),
    "g4_lts_c12": (dict(seed=1004, nchans=12, npts=18000, fs=100.0, aperture_km=0.8, flip_channel=3, shift_channel=1),
                   dict(nbands=2, fmin=0.5, fmax=5.0, alpha=0.5, winlen=("constant", 30.0), winover=0.5, order=2)),
    "g6_lts_c6_flip": (dict(seed=1006, nchans=6, npts=30000, fs=100.0, aperture_km=0.5, flip_channel=3),
                       dict(nbands=2, fmin=0.5, fmax=5.0, alpha=0.5, winlen=("constant", 30.0), winover=0.5, order=2)),
    "g5_ols_c5_overlap75": (dict(seed=1005, nchans=5, npts=20000, fs=50.0, aperture_km=0.3, noise_every_min=2),
                            dict(nbands=3, fmin=0.8, fmax=8.0, alpha=1.0, winlen=("constant", 20.0), winover=0.75, order=4)),
}


def run_case(name, ref):
    ref_helpers, ref_nbls, ref_cls, DataBin, ltsva = ref
    from narrow_band_least_squares_b200.synth import plane_wave_record
    skw, pkw = CASES[name]
    rec = plane_wave_record(**skw)
    st = make_stream(rec)
    nb = pkw["nbands"]
    freqlist, nb, _ = ref_helpers.get_freqlist(pkw["fmin"], pkw["fmax"], "log", nb)
    wl = pkw["winlen"]
    winlen_list = ref_helpers.get_winlenlist(wl[0], nb, wl[1], wl[1] if len(wl) < 3 else wl[1],
                                             wl[1] if len(wl) < 3 else wl[2])
    out = dict(input_sha=sha(rec.data), lat=np.array(rec.lat), lon=np.array(rec.lon), fs=rec.fs,
               freqlist=np.asarray(freqlist), winlen_list=np.asarray(winlen_list, dtype=np.float64),
               winover=pkw["winover"], alpha=pkw["alpha"], order=pkw["order"], nbands=nb)
    t0 = time.time()
    for ii in range(nb):
        stf, fs, sos = ref_helpers.filter_data(st, "butter", freqlist[ii], freqlist[ii + 1], pkw["order"], 0.01)
        filt = np.stack([tr.data for tr in stf])
        out[f"b{ii}_sos"] = sos
        out[f"b{ii}_filt_sha"] = sha(filt)
        out[f"b{ii}_filt_head"] = filt[:, :64].copy()
        out[f"b{ii}_filt_tail"] = filt[:, -64:].copy()
        out[f"b{ii}_filt_mid"] = filt[:, filt.shape[1] // 2: filt.shape[1] // 2 + 64].copy()
        out[f"b{ii}_filt_rms"] = np.sqrt(np.mean(filt ** 2, axis=1))
        res = run_estimator(ref_cls, DataBin, stf, rec.lat, rec.lon, winlen_list[ii], pkw["winover"], pkw["alpha"])
        for k, v in res.items():
            out[f"b{ii}_{k}"] = v
    # the public drivers, serial and joblib (narrow_band_least_squares.py:8,223)
    freq_resp_list = np.logspace(-2, np.log10(rec.fs / 2), 50)
    w0 = np.zeros(50)
    h0 = np.zeros(50, dtype=complex)
    args = (winlen_list, pkw["winover"], pkw["alpha"], st, list(rec.lat), list(rec.lon), nb, w0, h0, freqlist,
            "log", freq_resp_list, "butter", pkw["order"], 0.01)
    vel, baz, md, tt, stdict_all, sig, ncl, w_arr, h_arr = ref_nbls.narrow_band_least_squares(*args)
    # rows beyond num_compute_list are np.empty garbage in the serial driver: blank them
    for arr in (vel, baz, md, tt, sig):
        for ii in range(nb):
            arr[ii, ncl[ii]:] = np.nan
    if pkw["alpha"] != 1.0:
        sig[:] = np.nan      # never written for alpha<1 (narrow_band_least_squares.py:110-115)
    out.update(nbls_vel=vel, nbls_baz=baz, nbls_mdccm=md, nbls_t=tt, nbls_sig_tau=sig,
               nbls_num_compute=np.asarray(ncl), nbls_w=w_arr, nbls_h=h_arr, nbls_freq_resp_list=freq_resp_list)
    if stdict_all is not None:
        k, v, o, size = stdict_to_arrays(stdict_all)
        out.update(nbls_stdict_keys=k, nbls_stdict_vals=v, nbls_stdict_off=o, nbls_stdict_size=size)
    out["ref_seconds"] = time.time() - t0
    path = os.path.join(GOLD, name + ".npz")
    np.savez_compressed(path, **out)
    print(f"{name}: {os.path.getsize(path) / 1024:.0f} KiB in {out['ref_seconds']:.1f} s")


def run_tables(ref):
    """Known answers for the small pure functions (lts_classes.py:16-36,156-344,1013)."""
    _, _, ref_cls, _, _ = ref
    out = {}
    for P in (6, 10, 15, 28, 66, 300):
        seed = 0
        pairs = []
        seeds = []
        for _ in range(520):
            idx, seed = ref_cls.random_set(P, 2, seed)
            pairs.append(idx)
            seeds.append(seed)
        out[f"randset_{P}"] = np.array(pairs, dtype=np.int64)
        out[f"randseed_{P}"] = np.array(seeds, dtype=np.int64)
    rows = []
    for P in (6, 10, 15, 28, 66, 300):
        for alpha in (0.5, 0.6, 0.75, 0.875, 0.9, 0.99):
            nhalf = np.floor((P + 3) / 2)
            h = int(np.floor(2 * nhalf - P + 2 * (P - nhalf) * alpha))
            raw = ref_cls.raw_corfactor_lts(2, P, alpha) * ref_cls.raw_consfactor_lts(h, P)
            rew1 = ref_cls.rew_corfactor_lts(2, P, alpha)
            rows.append((P, alpha, h, raw, rew1))
            tab = np.full(P + 1, np.nan)
            with np.errstate(all="ignore"):
                for m in range(1, P):
                    w = np.zeros(P, dtype=bool)
                    w[:m] = True
                    tab[m] = ref_cls.rew_consfactor_lts(w, 2, P)
            out[f"rewcons_{P}"] = tab
    out["factors"] = np.array(rows)
    np.savez_compressed(os.path.join(GOLD, "tables.npz"), **out)
    print("tables written")


def run_helpers(ref):
    """Known answers for the host helpers on either side of the path: every band layout
    of get_freqlist (helpers.py:8-79), get_winlenlist (:83-104), the cheby1 branch of filter_data (:129-137) and
    the results text file (:161-235)."""
    import contextlib
    import io
    import tempfile
    ref_helpers = ref[0]
    from narrow_band_least_squares_b200.synth import plane_wave_record
    out = {}
    combos = [(0.5, 5.0, 8), (0.1, 10.0, 5), (1.0, 8.0, 3)]
    for kind in ("linear", "log", "octave", "2_octave_over", "onethird_octave", "octave_linear"):
        for ci, (fmin, fmax, nb) in enumerate(combos):
            try:
                fl, nbc, fmc = ref_helpers.get_freqlist(fmin, fmax, kind, nb)
            except ZeroDivisionError:        # octave_linear with no band left for the linear part (helpers.py:73)
                fl, nbc, fmc = [], -1, np.nan
            out[f"freq_{kind}_{ci}"] = np.asarray(fl, dtype=np.float64)
            out[f"freq_{kind}_{ci}_meta"] = np.array([fmin, fmax, nb, nbc, fmc], dtype=np.float64)
    for ci, (kind, nb, wl, w1, wx) in enumerate([("constant", 8, 30.0, 0, 0), ("adaptive", 20, 0, 60.0, 30.0),
                                                   ("adaptive", 7, 0, 45.0, 12.0)]):
        out[f"winlen_{ci}"] = np.asarray(ref_helpers.get_winlenlist(kind, nb, wl, w1, wx), dtype=np.float64)
        out[f"winlen_{ci}_meta"] = np.array([nb, wl, w1, wx], dtype=np.float64)
    rec = plane_wave_record(seed=1010, nchans=3, npts=9000, fs=50.0, aperture_km=0.3)
    st = make_stream(rec)
    for ci, (fmin, fmax, order, ripple) in enumerate([(1.0, 3.0, 2, 0.01), (0.5, 2.0, 4, 0.5)]):
        stf, fs, sos = ref_helpers.filter_data(st, "cheby1", fmin, fmax, order, ripple)
        filt = np.stack([tr.data for tr in stf])
        out[f"cheby_{ci}_meta"] = np.array([fmin, fmax, order, ripple], dtype=np.float64)
        out[f"cheby_{ci}_sos"] = sos
        out[f"cheby_{ci}_head"] = filt[:, :128].copy()
        out[f"cheby_{ci}_mid"] = filt[:, 4500:4628].copy()
        out[f"cheby_{ci}_tail"] = filt[:, -128:].copy()
        out[f"cheby_{ci}_rms"] = np.sqrt(np.mean(filt ** 2, axis=1))
    out["cheby_seed"] = np.array([1010, 3, 9000, 50.0, 0.3])
    # results text file written by the reference
    rng = np.random.default_rng(3)
    fl = np.array([0.5, 1.0, 2.0, 4.0])
    ncl = [4, 4, 5]          # read_txtfile sizes its arrays by the LAST band (helpers.py:216)
    arrs = [rng.random((3, 5)) for _ in range(4)]        # vel, baz, mdccm, t
    arrs[3] = 19000.0 + np.sort(arrs[3], axis=1)
    with tempfile.TemporaryDirectory() as d:
        with contextlib.redirect_stdout(io.StringIO()):
            ref_helpers.write_txtfile(d + "/", "ref", arrs[0], arrs[1], arrs[2], arrs[3], fl, ncl)
        out["txt_bytes"] = np.frombuffer(open(d + "/ref.txt", "rb").read(), dtype=np.uint8)
        back = ref_helpers.read_txtfile(d + "/", "ref")
    for k, v in zip(("vel", "baz", "mdccm", "t"), arrs):
        out[f"txt_{k}"] = v
    out["txt_freqlist"], out["txt_ncl"] = fl, np.asarray(ncl)
    for k, v in zip(("vel", "baz", "mdccm", "t", "freqlist", "ncl"), back[:6]):
        v = np.asarray(v, dtype=np.float64).copy()
        if v.ndim == 2:
            for ii in range(3):
                v[ii, ncl[ii]:] = np.nan          # np.empty garbage beyond a band's count
        out[f"txt_back_{k}"] = v
    out["txt_back_scalars"] = np.array([back[6], back[7], back[8]], dtype=np.float64)
    np.savez_compressed(os.path.join(GOLD, "helpers.npz"), **out)
    print("helpers written")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default=None)
    a = ap.parse_args()
    sys.path.insert(0, ROOT)
    ref = import_reference()
    os.makedirs(GOLD, exist_ok=True)
    if a.only in (None, "tables"):
        run_tables(ref)
    if a.only in (None, "helpers"):
        run_helpers(ref)
    for name in CASES:
        if a.only in (None, name):
            run_case(name, ref)


if __name__ == "__main__":
    main()
