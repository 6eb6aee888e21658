"""``ltsva`` with the reference's signature and return tuple, computed by the CUDA kernels.

Mirrors lts_array/lts_array/ltsva.py:9-54 and the estimator protocol of
lts_array/lts_array/classes/{lts_data_class.py:15-80, lts_classes.py:799-1036}:
``DataBin`` (window grid, tvec, co-array), ``correlate`` (K2), ``solve`` (K3).
The slowness-ellipse confidence intervals (lts_classes.py:347-640, :715-759, :938-968) are not on
the narrow-band path; ``conf_int_vel`` / ``conf_int_baz`` come from ``uncertainty.py``.
"""
from __future__ import annotations

import numpy as np

from . import _capi
from .engine import default_engine, window_grid
from .geometry import CoArray, get_rij
from .stream import as_record


class DataBin:
    """Window grid, time vector, data matrix and array coordinates — lts_data_class.py:15-80."""

    def __init__(self, window_length, window_overlap, alpha):
        self.window_length = window_length
        self.window_overlap = window_overlap
        self.alpha = alpha

    def build_data_arrays(self, st, latlist, lonlist, remove_elements=None, rij=None):
        if not isinstance(st, tuple) and hasattr(st, "__len__") and hasattr(st, "__getitem__") \
                and not hasattr(st, "fs"):
            if len(set([len(tr) for tr in st])) != 1:
                raise ValueError('Traces in stream must have same length!')
            if (remove_elements is not None) and (len(remove_elements) > 0):
                # like the reference, this edits the caller's st / lat / lon in place (:34-40)
                rem = np.sort(remove_elements)
                for jj in range(len(rem)):
                    st.remove(st[rem[jj]])
                    latlist.remove(latlist[rem[jj]])
                    lonlist.remove(lonlist[rem[jj]])
                    rem -= 1
            self.station_name = st[0].stats.station
            self.element_names = [tr.stats.station[-2:] if tr.stats.location == '' else tr.stats.location
                                  for tr in st]
        data, fs, tvec = as_record(st)
        self.nchans, self.npts = data.shape
        self.sampling_rate = fs
        self.winlensamp, self.sampinc, self.nits = window_grid(self.npts, fs, self.window_length,
                                                               self.window_overlap)
        self.intervals = np.arange(self.nits, dtype="int") * self.sampinc
        self.tvec = tvec if tvec is not None else np.arange(self.npts) / fs / 86400.0
        self.data_ct = data                      # channel-major (C, T): the device layout
        self.rij = np.asarray(rij, dtype=np.float64) if rij is not None else get_rij(latlist, lonlist, self.nchans)
        if np.shape(self.rij)[1] < 3:
            raise RuntimeError('The array must have at least 3 elements for well-posed least squares '
                               'estimation. Check rij array coordinates.')

    @property
    def data(self):
        """(npts, nchans) view, the reference's orientation (lts_data_class.py:68-70)."""
        return self.data_ct.T


class LsBeam:
    """Co-array + correlate — lts_classes.py:799-880."""

    def __init__(self, data: DataBin, engine=None, max_lag=-1, method=_capi.XCORR_AUTO):
        self.engine = engine if engine is not None else default_engine()
        self.dimension_number = 2
        n = data.nits
        self.mdccm = np.full(n, np.nan)
        self.t = np.full(n, np.nan)
        self.lts_vel = np.full(n, np.nan)
        self.lts_baz = np.full(n, np.nan)
        self.co = CoArray(data.rij)
        self.idx_pair = self.co.idx_pair
        self.xij, self.xij_mad, self.xij_standardized = self.co.xij, self.co.xij_mad, self.co.xij_standardized
        self.co_array_num = self.co.co_array_num
        self.tau = np.empty((self.co_array_num, n))
        self.time_delay_mad = np.zeros(n)
        self.conf_int_vel = np.full(n, np.nan)
        self.conf_int_baz = np.full(n, np.nan)
        self.sigma_tau = np.full(n, np.nan)
        self.stdict = {}
        self.p = 0.90
        self.max_lag, self.method = max_lag, method
        self.co.check_rank()

    def correlate(self, data: DataBin):
        n = data.nits
        half = int(data.winlensamp / 2)
        self.t = np.asarray(data.tvec)[data.intervals + half].astype(np.float64)      # :860
        lag, cmax = self.engine.xcorr(data.data_ct[None], [data.winlensamp], [max(data.sampinc, 1)], [n],
                                      max_lag=self.max_lag, method=self.method)
        self.lag = lag.T.copy()                                   # (P, W) integer lags
        self.cmax = cmax.T.copy()
        self.tau = (-self.lag.astype(np.float64) / data.sampling_rate).reshape(self.co_array_num, n, 1)
        self._lag_ew, self._cmax_ew = lag, cmax

    def _solve(self, data: DataBin, alpha):
        out = self.engine.solve(self._lag_ew, self._cmax_ew, self.co, data.sampling_rate, alpha)
        self.mdccm = out["mdccm"]
        self.time_delay_mad = out["mad"]
        self.lts_vel, self.lts_baz, self.sigma_tau = out["vel"], out["baz"], out["sigma_tau"]
        self.slowness = out["slowness"].T
        # slowness-ellipse uncertainties (lts_classes.py:715-759, :938-968): one vectorised pass over the windows
        from .uncertainty import slowness_ellipse_ci
        self.conf_int_vel, self.conf_int_baz = slowness_ellipse_ci(
            out["slowness"], self.sigma_tau, self.xij, out.get("weights") if alpha != 1.0 else None, self.p)
        return out


class OLSEstimator(LsBeam):
    """lts_classes.py:883-968."""

    def solve(self, data: DataBin):
        if data.nits:
            self._solve(data, 1.0)


class LTSEstimator(LsBeam):
    """lts_classes.py:971-1036."""

    def __init__(self, data: DataBin, **kw):
        super().__init__(data, **kw)
        self.element_weights = np.zeros((self.co_array_num, data.nits))
        if data.alpha == 1.0:
            raise RuntimeError('ALPHA = 1.0. This class is computionally inefficient. Use the OLSEstimator '
                               'class instead.')
        if np.shape(self.xij)[0] < (2 * self.dimension_number):
            raise RuntimeError('The co-array must have at least 4 elements for least trimmed squares. '
                               'Check rij array coordinates.')
        from .lts_tables import h_value
        self.h = h_value(self.co_array_num, data.alpha)

    def solve(self, data: DataBin):
        if data.nits:
            out = self._solve(data, data.alpha)
            self.element_weights = out["weights"].T.astype(np.float64)
        self.stdict = build_stdict(self.element_weights, self.t, self.co.idx, data.nchans)


def build_stdict(element_weights, t, idx, nchans):
    """Flagged element pairs per window — lts_classes.py:767-793, :1028-1036.

    ``idx`` is the (P, 2) pair table; values are 1-based element numbers, firsts then seconds.
    """
    stdict = {}
    nits = element_weights.shape[1]
    first, second = idx[:, 0] + 1, idx[:, 1] + 1
    for jj in range(nits):
        bad = np.where(element_weights[:, jj] == 0)[0]
        if len(bad) > 0:
            stdict[str(t[jj])] = np.concatenate((first[bad], second[bad]))
        if jj == nits - 1:
            stdict['size'] = nchans
    return stdict


def ltsva(st, lat_list, lon_list, window_length, window_overlap, alpha=1.0, plot_array_coordinates=False,
          remove_elements=None, *, rij=None, engine=None, max_lag=-1, method=_capi.XCORR_AUTO):
    """Least (trimmed) squares array processing of one (already filtered) record.

    Same positional arguments and 8-tuple as the reference's ``ltsva`` (ltsva.py:9,54):
    ``(lts_vel, lts_baz, t, mdccm, stdict, sigma_tau, conf_int_vel, conf_int_baz)``.
    Keyword-only extras: ``rij`` (2, C) km coordinates instead of lat/lon, ``engine``,
    ``max_lag`` (restrict the lag search; default = all 2N-1 lags like the reference), ``method``.
    ``plot_array_coordinates`` is accepted and ignored (plotting is out of scope).
    """
    data = DataBin(window_length, window_overlap, alpha)
    data.build_data_arrays(st, lat_list, lon_list, remove_elements, rij=rij)
    kw = dict(engine=engine, max_lag=max_lag, method=method)
    est = OLSEstimator(data, **kw) if data.alpha == 1.0 else LTSEstimator(data, **kw)
    est.correlate(data)
    est.solve(data)
    return (est.lts_vel, est.lts_baz, est.t, est.mdccm, est.stdict, est.sigma_tau, est.conf_int_vel,
            est.conf_int_baz)
