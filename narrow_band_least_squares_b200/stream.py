"""A minimal Stream/Trace pair and the adapter that turns any stream-like into a (C, T) record.

The reference consumes obspy ``Stream`` objects but touches only a small duck type
(SURVEY.md §8(b)): ``len(st)``, iteration/indexing, ``tr.data``, ``tr.stats.sampling_rate/npts/
station/location``, ``st[0].times('matplotlib')``, ``st.copy()``, ``st.remove(tr)``.  obspy is not a
dependency here: real obspy Streams work through the same duck type, and these two classes stand
in when obspy is absent.  (lts_data_class.py:30-70, helpers.py:124-139.)
"""
from __future__ import annotations

import copy

import numpy as np

MPL_EPOCH_OFFSET_DAYS = 0.0   # times('matplotlib') of these classes = starttime[s]/86400 + t/86400


class Stats(dict):
    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v


class Trace:
    def __init__(self, data, header=None):
        self.data = np.asarray(data)
        h = dict(sampling_rate=1.0, station="", location="", starttime=0.0)
        h.update(header or {})
        self.stats = Stats(h)
        self.stats["npts"] = len(self.data)

    def __len__(self):
        return len(self.data)

    def times(self, type="relative"):
        t = np.arange(len(self.data)) / float(self.stats.sampling_rate)
        if type == "matplotlib":
            return (float(self.stats.starttime) + t) / 86400.0 + MPL_EPOCH_OFFSET_DAYS
        return t

    def copy(self):
        return copy.deepcopy(self)


class Stream:
    def __init__(self, traces=None):
        self.traces = list(traces or [])

    def __len__(self):
        return len(self.traces)

    def __iter__(self):
        return iter(self.traces)

    def __getitem__(self, i):
        return self.traces[i]

    def append(self, tr):
        self.traces.append(tr)
        return self

    def remove(self, tr):
        self.traces.remove(tr)
        return self

    def copy(self):
        return copy.deepcopy(self)


def stream_from_array(data, fs, starttime=0.0, stations=None):
    data = np.asarray(data)
    return Stream([Trace(data[i], dict(sampling_rate=float(fs), starttime=starttime,
                                       station=(stations[i] if stations else f"E{i + 1:02d}"),
                                       location=f"{i + 1:02d}")) for i in range(data.shape[0])])


def as_record(st):
    """stream-like -> (data (C, T) float64 C-contiguous, fs, tvec or None).

    Accepts an obspy-like Stream, or a ``(data, fs[, tvec])`` tuple with data shaped (C, T).
    Mirrors the checks of DataBin.build_data_arrays (lts_data_class.py:30-31).
    """
    if isinstance(st, tuple):
        data = np.ascontiguousarray(st[0], dtype=np.float64)
        return data, float(st[1]), (np.asarray(st[2], dtype=np.float64) if len(st) > 2 and st[2] is not None else None)
    if hasattr(st, "data") and hasattr(st, "fs") and not hasattr(st, "stats"):   # SynthRecord-like
        tvec = (getattr(st, "starttime", 0.0) + np.arange(st.data.shape[1]) / st.fs) / 86400.0
        return np.ascontiguousarray(st.data, dtype=np.float64), float(st.fs), tvec
    n = len(st)
    if n == 0:
        raise ValueError("empty stream")
    lens = [len(tr) for tr in st]
    if any(m != lens[0] for m in lens):
        raise ValueError('Traces in stream must have same length!')
    data = np.empty((n, lens[0]), dtype=np.float64)
    for i, tr in enumerate(st):
        data[i] = tr.data
    fs = float(st[0].stats.sampling_rate)
    tvec = np.asarray(st[0].times('matplotlib'), dtype=np.float64)
    return data, fs, tvec


def like_stream(st, data):
    """A copy of ``st`` whose traces carry the rows of ``data`` (what filter_data returns)."""
    if isinstance(st, tuple):
        return (data,) + tuple(st[1:])
    if hasattr(st, "data") and hasattr(st, "fs") and not hasattr(st, "stats"):
        out = copy.copy(st)
        out.data = data
        return out
    out = st.copy()
    for i, tr in enumerate(out):
        tr.data = data[i]
    return out
