"""BASELINE.json workloads as job descriptions (SURVEY.md §8(a) header, §8(d)).

``config_job(name, hours=...)`` returns the host-side pieces of a run: band list, window lengths and
the processing constants, mirroring the constants block of the reference's example.py:41-72.
"""
from __future__ import annotations

import numpy as np

from .helpers import get_freqlist, get_winlenlist
from .synth import CONFIGS

FILTER_TYPE, FILTER_ORDER, FILTER_RIPPLE, WINOVER = "butter", 2, 0.01, 0.5


def config_job(name):
    c = CONFIGS[name]
    freqlist, nbands, _ = get_freqlist(c["fmin"], c["fmax"], "log", c["nbands"])
    wl = c["winlen"]
    if wl[0] == "constant":
        winlen_list = get_winlenlist("constant", nbands, wl[1], wl[1], wl[1])
    else:
        winlen_list = get_winlenlist("adaptive", nbands, wl[1], wl[1], wl[2])
    return dict(name=name, nchans=c["nchans"], fs=c["fs"], hours=c["hours"], nbands=nbands,
                freqlist=np.asarray(freqlist), winlen_list=list(winlen_list), winover=WINOVER, alpha=c["alpha"],
                filter_type=FILTER_TYPE, filter_order=FILTER_ORDER, filter_ripple=FILTER_RIPPLE,
                aperture_km=c["aperture_km"], seed=c["seed"])


def config_plan(name, npts, rij, **kw):
    """Plan of config ``name`` for a record of ``npts`` samples with element coordinates ``rij``."""
    from .narrow_band_least_squares import build_plan
    j = config_job(name)
    return build_plan(j["nchans"], npts, j["fs"], rij, j["winlen_list"], j["winover"], j["alpha"], j["nbands"],
                      j["freqlist"], "log", j["filter_type"], j["filter_order"], j["filter_ripple"], quiet=True, **kw)
