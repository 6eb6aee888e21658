"""The narrow-band least-squares drivers with the reference's positional signature and 9-tuple.

Mirrors narrow_band_least_squares.py of amiezzi/narrow_band_least_squares:
  narrow_band_least_squares           :8-127    (serial band loop)
  narrow_band_least_squares_parallel  :223-323  (joblib twin, same arguments and outputs)
Here every band and window goes to the GPU in ONE job (filter -> xcorr -> solve, libnbls'
nbls_run_host); both names run the same path — the "parallel" twin exists so callers of
example_parallel.py:138 can switch imports without edits.

Differences kept on purpose:
  * rows beyond ``num_compute_list[ii]`` are NaN (the reference leaves np.empty garbage / zeros);
  * ``sig_tau_array`` is filled for ALPHA < 1 too (the reference leaves it unset there);
  * if a band has more windows than the reference's ``vector_len`` sizing quirk allows (the
    reference then raises a broadcast error), the arrays are widened instead.
"""
from __future__ import annotations

import numpy as np
from scipy import signal

from . import _capi
from .engine import Plan, Results, default_engine, window_grid
from .geometry import CoArray, get_rij
from .helpers import band_edges, design_sos, plot_sos
from .ltsva import build_stdict
from .stream import as_record


def reference_vector_len(npts, Fs, WINLEN_list, WINOVER):
    """The reference's output width — narrow_band_least_squares.py:41-47 (uses the LAST band's
    window length in seconds as a sample step, then divides by Fs)."""
    sampinc = int((1 - WINOVER) * WINLEN_list[-1])
    nits = len(np.arange(0, npts, sampinc)) - 1
    return int(nits / Fs)


def build_plan(nchan, npts, Fs, rij, WINLEN_list, WINOVER, ALPHA, NBANDS, freqlist, FREQ_BAND_TYPE, FILTER_TYPE,
               FILTER_ORDER, FILTER_RIPPLE, max_lag=-1, method=_capi.XCORR_AUTO, quiet=False, **plan_kw):
    """Host-side job description for all bands: filter sections, window grids, co-array, LTS tables."""
    sos_list, n_l, h_l, w_l = [], [], [], []
    for ii in range(NBANDS):
        fmin, fmax = band_edges(freqlist, ii, FREQ_BAND_TYPE)
        sos_list.append(design_sos(FILTER_TYPE, fmin, fmax, Fs, FILTER_ORDER, FILTER_RIPPLE))
        bt = WINLEN_list[ii] * (fmax - fmin)
        if bt < 5.0 and not quiet:     # narrow_band_least_squares.py:83-87
            print('CAUTION: BT < 5! Band between ' + str(fmin) + ' Hz and ' + str(fmax) + ' Hz has BT = ' + str(bt))
        n, hop, w = window_grid(npts, Fs, WINLEN_list[ii], WINOVER)
        n_l.append(n); h_l.append(max(hop, 1)); w_l.append(w)
    nsec = {s.shape[0] for s in sos_list}
    if len(nsec) != 1:
        raise ValueError("all bands must have the same number of filter sections")
    return Plan(nchan, npts, Fs, np.stack(sos_list), FILTER_TYPE == 'butter', n_l, h_l, w_l, rij, alpha=ALPHA,
                max_lag=max_lag, method=method, **plan_kw)


def prepare_job(WINLEN_list, WINOVER, ALPHA, st, lat_list, lon_list, NBANDS, freqlist, FREQ_BAND_TYPE, FILTER_TYPE,
                FILTER_ORDER, FILTER_RIPPLE, rij=None, max_lag=-1, method=_capi.XCORR_AUTO):
    """Host side of one driver call before the GPU runs: samples, time vector, co-array checks, job description."""
    data, Fs, tvec = as_record(st)
    nchan, npts = data.shape
    if rij is None:
        rij = get_rij(lat_list, lon_list, nchan)
    if np.shape(rij)[1] < 3:
        raise RuntimeError('The array must have at least 3 elements for well-posed least squares estimation. '
                           'Check rij array coordinates.')
    co = CoArray(np.asarray(rij, dtype=np.float64))
    co.check_rank()
    if ALPHA != 1.0 and co.co_array_num < 4:
        raise RuntimeError('The co-array must have at least 4 elements for least trimmed squares. '
                           'Check rij array coordinates.')
    plan = build_plan(nchan, npts, Fs, co, WINLEN_list, WINOVER, ALPHA, NBANDS, freqlist, FREQ_BAND_TYPE,
                      FILTER_TYPE, FILTER_ORDER, FILTER_RIPPLE, max_lag=max_lag, method=method)
    if tvec is None:
        tvec = np.arange(npts) / Fs / 86400.0
    return data, tvec, co, plan


def assemble_outputs(plan, res, tvec, co, WINLEN_list, WINOVER, ALPHA, NBANDS, w, h, freqlist, FREQ_BAND_TYPE,
                     freq_resp_list, FILTER_TYPE, FILTER_ORDER, FILTER_RIPPLE):
    """Results of one GPU job -> the reference's 9-tuple (narrow_band_least_squares.py:94-127)."""
    Fs, npts, nchan = plan.fs, plan.npts, plan.nchan
    vector_len = max(reference_vector_len(npts, Fs, WINLEN_list, WINOVER), int(plan.win_count.max(initial=0)))
    vel_array, baz_array, mdccm_array, sig_tau_array, t_array = (np.full((NBANDS, vector_len), np.nan)
                                                                 for _ in range(5))
    w_array = np.empty((NBANDS, len(w)), dtype=complex)
    h_array = np.empty((NBANDS, len(h)), dtype=complex)
    stdict_all = None if ALPHA == 1.0 else {}
    num_compute_list = []
    for ii in range(NBANDS):
        fmin, fmax = band_edges(freqlist, ii, FREQ_BAND_TYPE)
        sos = plot_sos(FILTER_TYPE, fmin, fmax, Fs, FILTER_ORDER, FILTER_RIPPLE)
        w_array[ii, :], h_array[ii, :] = signal.sosfreqz(sos, freq_resp_list, fs=Fs)
        r = plan.band_rows(ii)
        nw = int(plan.win_count[ii])
        t = tvec[np.arange(nw) * int(plan.win_hop[ii]) + int(int(plan.win_len[ii]) / 2)]   # lts_classes.py:860
        vel_array[ii, :nw] = res.vel[r]
        baz_array[ii, :nw] = res.baz[r]
        mdccm_array[ii, :nw] = res.mdccm[r]
        sig_tau_array[ii, :nw] = res.sigma_tau[r]
        t_array[ii, :nw] = t
        num_compute_list.append(nw)
        if ALPHA != 1.0:
            sd = build_stdict(res.weights[r].T, t, co.idx, nchan)
            for key, val in sd.items():       # narrow_band_least_squares.py:116-124
                stdict_all[key if key == 'size' else str(ii + 1).zfill(2) + '_' + key] = val
    return (vel_array, baz_array, mdccm_array, t_array, stdict_all, sig_tau_array, num_compute_list, w_array, h_array)


def narrow_band_least_squares(WINLEN_list, WINOVER, ALPHA, st, lat_list, lon_list, NBANDS, w, h, freqlist,
                              FREQ_BAND_TYPE, freq_resp_list, FILTER_TYPE, FILTER_ORDER, FILTER_RIPPLE, *,
                              rij=None, engine=None, max_lag=-1, method=_capi.XCORR_AUTO, return_details=False):
    """Narrow-band least-squares array processing of ``st`` in ``NBANDS`` frequency bands.

    Returns ``(vel_array, baz_array, mdccm_array, t_array, stdict_all, sig_tau_array,
    num_compute_list, w_array, h_array)`` exactly like the reference (arrays shaped
    (NBANDS, vector_len); ``stdict_all`` is None for ALPHA == 1.0, else keys "BB_<t>" + 'size').
    """
    data, tvec, co, plan = prepare_job(WINLEN_list, WINOVER, ALPHA, st, lat_list, lon_list, NBANDS, freqlist,
                                       FREQ_BAND_TYPE, FILTER_TYPE, FILTER_ORDER, FILTER_RIPPLE, rij=rij,
                                       max_lag=max_lag, method=method)
    eng = engine if engine is not None else default_engine()
    res = eng.run(plan, data, Results(plan, pinned=False, want_cmax=return_details))
    out = assemble_outputs(plan, res, tvec, co, WINLEN_list, WINOVER, ALPHA, NBANDS, w, h, freqlist, FREQ_BAND_TYPE,
                           freq_resp_list, FILTER_TYPE, FILTER_ORDER, FILTER_RIPPLE)
    if return_details:
        return out, plan, res
    return out


def narrow_band_least_squares_parallel(*args, **kwargs):
    """Same arguments and outputs as ``narrow_band_least_squares`` (reference :223-323)."""
    return narrow_band_least_squares(*args, **kwargs)
