"""Locate, stage and import the UNMODIFIED reference so that it can be executed as a checker / CPU baseline.

TEST INFRASTRUCTURE ONLY (tests/, __graft_entry__, bench.py's CPU legs, oracle/run_reference.py).  Nothing in the
product package imports this module.

The reference (amiezzi/narrow_band_least_squares) is pure Python, so "building" it for the GPU box means putting
the handful of files its hot path imports where the box can see them: ``stage_reference()`` — called from
``__graft_entry__.build()`` whenever /root/reference is present — packs them into the git-ignored archive
``oracle/_ref/reference_hotpath.zip`` (a build artefact: never in the tracked tree, it travels with gpurun like a built
.so and is executed straight from the archive with zipimport).  ``import_reference()`` then imports the reference
unmodified through the import shims of ``oracle/refstubs`` (obspy / matplotlib stand-ins, SURVEY.md
Appendix A.5) and replaces the module-local ``np`` of narrow_band_least_squares.py by a proxy that maps the
``'complex_'`` dtype string numpy >= 2 removed (narrow_band_least_squares.py:58-59,276-277) to complex128.
"""
from __future__ import annotations

import os
import shutil
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
STAGED = os.path.join(HERE, "_ref")
ARCHIVE = os.path.join(STAGED, "reference_hotpath.zip")      # the build artefact: imported with zipimport, never unpacked
# files on the hot path; plotting, waveform download and docs are not staged
HOT_PATH_FILES = (
    "narrow_band_least_squares.py",
    "helpers.py",
    "lts_array/lts_array/__init__.py",
    "lts_array/lts_array/ltsva.py",
    "lts_array/lts_array/classes/__init__.py",
    "lts_array/lts_array/classes/lts_classes.py",
    "lts_array/lts_array/classes/lts_data_class.py",
    "lts_array/lts_array/tools/__init__.py",
    "lts_array/lts_array/tools/lts_array_plot.py",
)


def reference_dir():
    """The reference to execute: $NBLS_REFERENCE, /root/reference (build container), else the staged archive
    oracle/_ref/reference_hotpath.zip (a path Python's zipimport accepts like a directory)."""
    for d in (os.environ.get("NBLS_REFERENCE"), "/root/reference"):
        if d and os.path.isfile(os.path.join(d, "narrow_band_least_squares.py")):
            return d
    if os.path.isfile(ARCHIVE):
        return ARCHIVE
    return None


def stage_reference(src="/root/reference", dst=ARCHIVE):
    """Pack the reference's hot-path files from ``src`` into the git-ignored archive ``dst`` (the "built" reference that
    travels to the GPU box; it is executed from the archive, unmodified, and never enters the tracked tree).  Returns the
    number of files packed (0 if ``src`` does not exist, e.g. on the GPU box, where the archive arrives with the snapshot)."""
    import zipfile
    if not os.path.isfile(os.path.join(src, "narrow_band_least_squares.py")):
        return 0
    os.makedirs(os.path.dirname(dst), exist_ok=True)
    for stale in HOT_PATH_FILES:                     # loose copies of an earlier layout
        f = os.path.join(os.path.dirname(dst), stale)
        if os.path.isfile(f):
            os.remove(f)
    shutil.rmtree(os.path.join(os.path.dirname(dst), "lts_array"), ignore_errors=True)
    n = 0
    tmp = dst + ".tmp"
    with zipfile.ZipFile(tmp, "w", zipfile.ZIP_DEFLATED) as z:
        for rel in HOT_PATH_FILES:
            s = os.path.join(src, rel)
            if os.path.isfile(s):
                z.write(s, rel)
                n += 1
    os.replace(tmp, dst)
    return n


class _NP:
    """numpy proxy: only 'complex_' dtype strings are rewritten."""

    def __getattr__(self, k):
        return getattr(np, k)

    @staticmethod
    def _fix(kw):
        if kw.get("dtype") == "complex_":
            kw["dtype"] = np.complex128
        return kw

    def empty(self, *a, **k):
        return np.empty(*a, **self._fix(k))

    def zeros(self, *a, **k):
        return np.zeros(*a, **self._fix(k))


def import_reference(ref=None):
    """-> (helpers, narrow_band_least_squares, lts_classes, DataBin, ltsva) modules/objects of the reference."""
    ref = ref or reference_dir()
    if ref is None:
        raise RuntimeError("reference tree not found (neither /root/reference nor the staged oracle/_ref)")
    stubs = os.path.join(HERE, "refstubs")
    paths = [stubs, ref, os.path.join(ref, "lts_array")]
    for p in reversed(paths):
        if p in sys.path:
            sys.path.remove(p)
        sys.path.insert(0, p)
    # joblib/loky workers re-import the reference's modules: they inherit PYTHONPATH
    old = [p for p in os.environ.get("PYTHONPATH", "").split(os.pathsep) if p and p not in paths]
    os.environ["PYTHONPATH"] = os.pathsep.join(paths + old)
    import lts_array  # noqa: F401
    import helpers as ref_helpers
    import narrow_band_least_squares as ref_nbls
    from lts_array.classes import lts_classes as ref_cls
    from lts_array.classes.lts_data_class import DataBin

    ref_nbls.np = _NP()
    return ref_helpers, ref_nbls, ref_cls, DataBin, lts_array.ltsva


def make_stream(rec):
    """SynthRecord -> the Stream duck type the reference consumes (refstubs' obspy stand-in)."""
    stubs = os.path.join(HERE, "refstubs")
    if stubs not in sys.path:
        sys.path.insert(0, stubs)
    from obspy import Stream, Trace
    st = Stream()
    for i in range(rec.nchans):
        st.append(Trace(rec.data[i].copy(), dict(sampling_rate=rec.fs, station=f"SYN{i + 1:02d}",
                                                 location=f"{i + 1:02d}", starttime=rec.starttime)))
    return st


def run_reference_driver(ref, rec, freqlist, winlen_list, winover, alpha, nbands, ftype="butter", order=2,
                         ripple=0.01, parallel=True):
    """Call the reference's own narrow_band_least_squares(_parallel) (narrow_band_least_squares.py:8 / :223) with the
    positional arguments example.py:134-140 / example_parallel.py:134-138 pass.  Returns its 9-tuple."""
    _, ref_nbls, _, _, _ = ref
    st = make_stream(rec)
    freq_resp_list = np.logspace(-2, np.log10(rec.fs / 2), 50)
    w0 = np.zeros(50)
    h0 = np.zeros(50, dtype=complex)
    fn = ref_nbls.narrow_band_least_squares_parallel if parallel else ref_nbls.narrow_band_least_squares
    return fn(list(winlen_list), winover, alpha, st, list(rec.lat), list(rec.lon), nbands, w0, h0, freqlist, "log",
              freq_resp_list, ftype, order, ripple)
