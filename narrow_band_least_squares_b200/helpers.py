"""Host-side configuration helpers with the reference's names and argument meaning (helpers.py of
amiezzi/narrow_band_least_squares): band lists, window-length lists, filter design, taper ramps,
result text files.  No sample arithmetic happens here — ``filter_data`` hands the record to the CUDA
filter kernel through libnbls.

Reference lines: get_freqlist helpers.py:8-79, get_winlenlist :83-104, filter_data :108-141,
make_float :145-158, write_txtfile/read_txtfile :161-235, get_rij :239-284.
"""
from __future__ import annotations

import math

import numpy as np
from scipy import signal

from .geometry import get_rij  # noqa: F401  (re-exported like the reference's helpers.get_rij)


def get_freqlist(FMIN, FMAX, FREQ_BAND_TYPE, NBANDS):
    """Narrow-band limits; returns (freqlist, nbands_calc, FMAX_calc) like helpers.py:8-79."""
    if FREQ_BAND_TYPE == 'linear':
        step = (FMAX - FMIN) / NBANDS
        return np.arange(FMIN, FMAX + step, step), NBANDS, FMAX
    if FREQ_BAND_TYPE == 'log':
        return np.logspace(math.log(FMIN, 10), math.log(FMAX, 10), num=NBANDS + 1), NBANDS, FMAX
    if FREQ_BAND_TYPE in ('octave', '2_octave_over'):
        fl = [FMIN]
        while 2 * fl[-1] <= FMAX:
            fl.append(2 * fl[-1])
        return fl, int(len(fl)) - (1 if FREQ_BAND_TYPE == 'octave' else 2), fl[-1]
    if FREQ_BAND_TYPE == 'onethird_octave':
        fl = [FMIN]
        while fl[-1] * (2 ** (1. / 3.)) <= FMAX:
            fl.append(fl[-1] * (2 ** (1. / 3.)))
        return fl, int(len(fl)) - 1, fl[-1]
    if FREQ_BAND_TYPE == 'octave_linear':
        switch_freq = 2
        fl = [FMIN]
        while 2 * fl[-1] <= switch_freq:
            fl.append(2 * fl[-1])
        rest = NBANDS - len(fl)
        step = (FMAX - fl[-1]) / rest
        fl = fl + list(np.arange(fl[-1], FMAX + step, step))
        return fl, int(len(fl)) - 1, FMAX
    raise ValueError(f"unknown FREQ_BAND_TYPE {FREQ_BAND_TYPE!r}")


def get_winlenlist(WINDOW_LENGTH_TYPE, NBANDS, WINLEN, WINLEN_1, WINLEN_X):
    """Window length [s] per band: 'constant' or 'adaptive' (linear, truncated to int) — helpers.py:83-104."""
    if WINDOW_LENGTH_TYPE == 'constant':
        return [WINLEN for _ in range(NBANDS)]
    if WINDOW_LENGTH_TYPE == 'adaptive':
        return [int(v) for v in np.linspace(WINLEN_1, WINLEN_X, num=NBANDS)]
    raise ValueError(f"unknown WINDOW_LENGTH_TYPE {WINDOW_LENGTH_TYPE!r}")


def band_edges(freqlist, ii, FREQ_BAND_TYPE):
    """(fmin, fmax) of band ii — narrow_band_least_squares.py:69-75."""
    if FREQ_BAND_TYPE == '2_octave_over':
        return freqlist[ii], freqlist[ii + 2]
    return freqlist[ii], freqlist[ii + 1]


def design_sos(FILTER_TYPE, FMIN, FMAX, Fs, FILTER_ORDER, FILTER_RIPPLE=None):
    """Second-order sections the filter kernel runs for one band.

    'butter': obspy's bandpass as called at helpers.py:127 (normalised corners -> zpk -> zpk2sos;
    falls back to a high-pass when fmax reaches Nyquist).  'cheby1': helpers.py:130.
    """
    if FILTER_TYPE == 'butter':
        fe = 0.5 * Fs
        low, high = FMIN / fe, FMAX / fe
        if high - 1.0 > -1e-6:
            import warnings
            warnings.warn(f"Selected high corner frequency ({FMAX}) of bandpass is at or above Nyquist ({fe}). "
                          "Applying a high-pass instead.")
            z, p, k = signal.iirfilter(FILTER_ORDER, low, btype='highpass', ftype='butter', output='zpk')
        else:
            if low > 1:
                raise ValueError("Selected low corner frequency is above Nyquist.")
            z, p, k = signal.iirfilter(FILTER_ORDER, [low, high], btype='band', ftype='butter', output='zpk')
        return signal.zpk2sos(z, p, k)
    if FILTER_TYPE == 'cheby1':
        return signal.iirfilter(FILTER_ORDER, [FMIN, FMAX], rp=FILTER_RIPPLE, btype='band', analog=False,
                                ftype='cheby1', fs=Fs, output='sos')
    raise ValueError(f"unknown FILTER_TYPE {FILTER_TYPE!r}")


def plot_sos(FILTER_TYPE, FMIN, FMAX, Fs, FILTER_ORDER, FILTER_RIPPLE=None):
    """The ``sos`` the reference's filter_data RETURNS (used only for sosfreqz plots): helpers.py:128,130."""
    if FILTER_TYPE == 'butter':
        return signal.iirfilter(FILTER_ORDER, [FMIN, FMAX], btype='band', ftype='butter', fs=Fs, output='sos')
    return design_sos(FILTER_TYPE, FMIN, FMAX, Fs, FILTER_ORDER, FILTER_RIPPLE)


def taper_ramps(npts, max_percentage=0.01):
    """(wlen, up, down): the Hann half-ramps of obspy Trace.taper(max_percentage) — helpers.py:139."""
    wlen = min(int(max_percentage * npts), int(npts / 2))
    if wlen <= 0:
        return 0, np.zeros(0), np.zeros(0)
    w = signal.windows.hann(2 * wlen if 2 * wlen == npts else 2 * wlen + 1)
    return wlen, np.ascontiguousarray(w[:wlen]), np.ascontiguousarray(w[len(w) - wlen:])


def make_float(input_list):
    """helpers.py:145-158: the values as a float ndarray."""
    return np.array([float(v) for v in input_list])


def filter_data(st, FILTER_TYPE, FMIN, FMAX, FILTER_ORDER, FILTER_RIPPLE):
    """Filter and taper one band on the GPU; returns (stf, Fs, sos) like helpers.py:108-141.

    ``st`` is any stream-like (see ``stream.Stream``); the input is not modified.
    """
    from .engine import default_engine
    from .stream import as_record, like_stream
    data, fs, _ = as_record(st)
    sos = design_sos(FILTER_TYPE, FMIN, FMAX, fs, FILTER_ORDER, FILTER_RIPPLE)
    filt = default_engine().filter_bands(data, fs, [sos], zerophase=(FILTER_TYPE == 'butter'))[0]
    return like_stream(st, filt), fs, plot_sos(FILTER_TYPE, FMIN, FMAX, fs, FILTER_ORDER, FILTER_RIPPLE)


def write_txtfile(save_dir, fname, vel_array, baz_array, mdccm_array, t_array, freqlist, num_compute_list):
    """Results text file, one row per band-window — same layout as helpers.py:159-181:
    header 'Fmin \t Fmax \t Time \t Trace_vel \t Backaz \t MdCCM', tab-separated str() values,
    written to ``save_dir + fname + '.txt'``."""
    with open(save_dir + fname + '.txt', 'w') as f:
        f.write('Fmin \t Fmax \t Time \t Trace_vel \t Backaz \t MdCCM \n')
        for ii in range(len(freqlist) - 1):
            for jj in range(num_compute_list[ii]):
                cols = (freqlist[ii], freqlist[ii + 1], t_array[ii, jj], vel_array[ii, jj], baz_array[ii, jj],
                        mdccm_array[ii, jj])
                f.write('\t'.join(str(c) for c in cols) + '\n')


def read_txtfile(save_dir, fname):
    """Inverse of write_txtfile — helpers.py:184-235.  Returns (vel_array, baz_array, mdccm_array, t_array,
    freqlist, num_compute_list, nbands, FMIN, FMAX); rows are padded to the LAST band's count like the
    reference (rows beyond a band's count are NaN here, uninitialised there)."""
    tab = np.genfromtxt(save_dir + fname + '.txt', skip_header=1, dtype='float')
    tab = np.atleast_2d(tab)
    fmin_col = tab[:, 0]
    uniq, idx = np.unique(fmin_col, return_index=True)
    freqlist = np.append(uniq, tab[-1, 1])
    idx = np.append(idx, len(fmin_col))
    num_compute_list = np.diff(idx)
    vector_len = len(fmin_col) - idx[-2]
    nbands = len(freqlist) - 1
    arrs = [np.full((nbands, vector_len), np.nan) for _ in range(4)]     # vel, baz, mdccm, t
    for ii in range(nbands):
        a, b = idx[ii], idx[ii + 1]
        n = min(b - a, vector_len)
        for arr, col in zip(arrs, (3, 4, 5, 2)):
            arr[ii, :n] = tab[a:a + n, col]
    return arrs[0], arrs[1], arrs[2], arrs[3], freqlist, num_compute_list, nbands, fmin_col[0], tab[-1, 1]
