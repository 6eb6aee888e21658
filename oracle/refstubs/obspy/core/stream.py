"""Stream/Trace duck types with obspy's documented band-pass and taper arithmetic (test infrastructure only).

obspy is not vendored by the reference and is not installable here; the
arithmetic below follows SURVEY.md Appendix A.1 (bandpass, zerophase) and
A.3 (Hann taper).  Parity at this boundary is unpinned.
"""
import copy as _copy
import warnings

import numpy as np
from scipy import signal


class UTCDateTime(float):
    """Seconds since the Unix epoch; just enough for Trace.times('matplotlib')."""


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
        st = Stats(sampling_rate=1.0, station="", location="", starttime=0.0,
                   network="", channel="")
        if header:
            st.update(header)
        st["npts"] = len(self.data)
        st["delta"] = 1.0 / st["sampling_rate"]
        self.stats = st

    def __len__(self):
        return len(self.data)

    def copy(self):
        return _copy.deepcopy(self)

    def times(self, type="relative"):
        rel = np.arange(self.stats.npts) / self.stats.sampling_rate
        if type == "relative":
            return rel
        if type == "matplotlib":
            # days since 0001-01-01 +1 in old matplotlib, days since 1970 in >=3.3; the
            # reference only carries the numbers through, so use days since 1970.
            return (float(self.stats.starttime) + rel) / 86400.0
        raise ValueError(type)

    def filter(self, type, **options):
        if type != "bandpass":
            raise NotImplementedError(type)
        fs = self.stats.sampling_rate
        self.data = _bandpass(self.data, options["freqmin"], options["freqmax"], fs,
                              corners=options.get("corners", 4),
                              zerophase=options.get("zerophase", False))
        return self

    def taper(self, max_percentage, type="hann", max_length=None, side="both"):
        npts = self.stats.npts
        wlen = min(int(max_percentage * npts), int(npts / 2)) if max_percentage is not None else int(npts / 2)
        if 2 * wlen == npts:
            w = signal.windows.hann(2 * wlen)
        else:
            w = signal.windows.hann(2 * wlen + 1)
        full = np.hstack((w[:wlen], np.ones(npts - 2 * wlen), w[len(w) - wlen:]))
        self.data = self.data * full
        return self


def _bandpass(data, freqmin, freqmax, df, corners=4, zerophase=False):
    fe = 0.5 * df
    low = freqmin / fe
    high = freqmax / fe
    if high - 1.0 > -1e-6:
        warnings.warn("freqmax at/above Nyquist: applying a high-pass instead")
        z, p, k = signal.iirfilter(corners, low, btype="highpass", ftype="butter", output="zpk")
    else:
        if low > 1:
            raise ValueError("Selected low corner frequency is above Nyquist.")
        z, p, k = signal.iirfilter(corners, [low, high], btype="band", ftype="butter", output="zpk")
    sos = signal.zpk2sos(z, p, k)
    if zerophase:
        firstpass = signal.sosfilt(sos, data)
        return signal.sosfilt(sos, firstpass[::-1])[::-1]
    return signal.sosfilt(sos, data)


class Stream:
    def __init__(self, traces=None):
        self.traces = list(traces) if traces is not None else []

    def __len__(self):
        return len(self.traces)

    def __iter__(self):
        return iter(self.traces)

    def __getitem__(self, i):
        if isinstance(i, slice):
            return Stream(self.traces[i])
        return self.traces[i]

    def copy(self):
        return _copy.deepcopy(self)

    def remove(self, tr):
        self.traces.remove(tr)
        return self

    def append(self, tr):
        self.traces.append(tr)
        return self

    def filter(self, type, **options):
        for tr in self.traces:
            tr.filter(type, **options)
        return self

    def taper(self, *a, **k):
        for tr in self.traces:
            tr.taper(*a, **k)
        return self
