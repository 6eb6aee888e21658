"""Engine: one libnbls context on one GPU, and the job description the C-ABI consumes.

All sample arithmetic runs in the CUDA kernels of ``csrc/`` through ``_capi`` (ctypes).  This
module only prepares small host-side tables (window grid, co-array, filter coefficients, LTS
tables) and moves buffers.  There is no CPU path: constructing an Engine without a CUDA device or
without the built library raises ``NBLSError``.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _capi
from ._capi import NBLSError, check
from .geometry import CoArray
from .helpers import taper_ramps
from . import lts_tables


def window_grid(npts, fs, window_length, window_overlap):
    """(N, hop, W): DataBin's grid — lts_data_class.py:58-64 (a window ending exactly at npts is excluded)."""
    n = int(window_length * fs)
    hop = int(np.round((1 - window_overlap) * n))
    if n <= 0:
        raise ValueError("window_length too short for the sampling rate")
    starts = np.arange(0, npts - n, hop, dtype="int") if hop > 0 else np.zeros(1 if npts > n else 0, dtype="int")
    return n, hop, len(starts)


class Plan:
    """Everything about a job that does not depend on the samples; keeps the host arrays alive."""

    def __init__(self, nchan, npts, fs, sos, zerophase, win_len, win_hop, win_count, rij, alpha=1.0,
                 win_first=None, max_lag=-1, method=_capi.XCORR_AUTO, taper=0.01, chunk_len=0, exact_scan=False,
                 global_t0=0, global_npts=None):
        self.nchan, self.npts, self.fs = int(nchan), int(npts), float(fs)
        self.sos = np.ascontiguousarray(sos, dtype=np.float64)
        if self.sos.ndim != 3 or self.sos.shape[2] != 6:
            raise ValueError("sos must have shape (nbands, nsections, 6)")
        self.nbands, self.nsections = self.sos.shape[:2]
        self.win_len = np.ascontiguousarray(win_len, dtype=np.int32)
        self.win_hop = np.ascontiguousarray(win_hop, dtype=np.int32)
        self.win_count = np.ascontiguousarray(win_count, dtype=np.int32)
        self.win_first = np.ascontiguousarray(win_first if win_first is not None else np.zeros(self.nbands),
                                              dtype=np.int64)
        for a in (self.win_len, self.win_hop, self.win_count, self.win_first):
            if a.shape != (self.nbands,):
                raise ValueError("window arrays must have one entry per band")
        self.offsets = np.concatenate(([0], np.cumsum(self.win_count))).astype(np.int64)
        self.nwin = int(self.offsets[-1])
        self.alpha = float(alpha)
        self.co = rij if isinstance(rij, CoArray) else CoArray(np.asarray(rij, dtype=np.float64))
        if self.co.nchans != self.nchan:
            raise ValueError('Mismatch between the number of stream channels and the array coordinates.')
        self.npairs = self.co.co_array_num
        g = global_npts if global_npts is not None else self.npts
        self.taper_len, self.taper_up, self.taper_down = taper_ramps(g, taper) if taper else (0, np.zeros(0), np.zeros(0))

        job = _capi.Job()
        f, w, s = job.filter, job.window, job.solve
        f.nbands, f.nsections, f.sos = self.nbands, self.nsections, _capi.dptr(self.sos)
        f.zerophase, f.exact_scan, f.chunk_len = int(bool(zerophase)), int(bool(exact_scan)), int(chunk_len)
        f.taper_len = self.taper_len
        f.taper_up, f.taper_down = _capi.dptr(self.taper_up), _capi.dptr(self.taper_down)
        f.global_t0, f.global_npts = int(global_t0), int(g)
        w.nbands, w.max_lag, w.method = self.nbands, int(max_lag), int(method)
        w.win_len, w.win_hop = _capi.dptr(self.win_len, _capi.c_int32_p), _capi.dptr(self.win_hop, _capi.c_int32_p)
        w.win_first, w.win_count = _capi.dptr(self.win_first, _capi.c_int64_p), _capi.dptr(self.win_count, _capi.c_int32_p)
        self._fill_solve(s)
        self.job = job

    def _fill_solve(self, s):
        co = self.co
        self.xij = np.ascontiguousarray(co.xij, dtype=np.float64)
        self.xij_mad = np.ascontiguousarray(co.xij_mad, dtype=np.float64)
        q, r = np.linalg.qr(co.xij_standardized)                      # lts_classes.py:889
        self.gain = np.ascontiguousarray(np.linalg.solve(r, q.T), dtype=np.float64)   # (2, P) = pinv(X_std)
        s.nchan, s.npairs, s.fs = self.nchan, self.npairs, self.fs
        s.xij, s.xij_mad, s.ols_gain = _capi.dptr(self.xij), _capi.dptr(self.xij_mad), _capi.dptr(self.gain)
        s.alpha = self.alpha
        if self.alpha != 1.0:
            h, raw, rew1, rew_tab = lts_tables.lts_constants(self.npairs, self.alpha)
            self.h = h
            self.subsets = np.ascontiguousarray(lts_tables.subset_stream(self.npairs), dtype=np.int32)
            self.rew_cons = np.ascontiguousarray(rew_tab, dtype=np.float64)
            s.h, s.n_starts, s.n_keep = h, lts_tables.N_STARTS, lts_tables.N_KEEP
            s.csteps1, s.csteps2 = lts_tables.CSTEPS1, lts_tables.CSTEPS2
            s.subsets = _capi.dptr(self.subsets, _capi.c_int32_p)
            s.raw_factor, s.rew_factor1 = raw, rew1
            s.rew_cons = _capi.dptr(self.rew_cons)

    def band_rows(self, b):
        return slice(int(self.offsets[b]), int(self.offsets[b + 1]))

    def with_alpha(self, alpha):
        """The same job with another solver (OLS <-> FAST-LTS): same filter, window grid and co-array."""
        import copy
        p = copy.copy(self)
        p.alpha = float(alpha)
        job = _capi.Job()
        C.memmove(C.byref(job), C.byref(self.job), C.sizeof(_capi.Job))
        p.job = job
        p._fill_solve(job.solve)
        return p


class Results:
    """Host arrays of one run, window-major over all bands (row e = offsets[b] + w)."""

    def __init__(self, plan: Plan, pinned=True, want_cmax=True):
        mk = _capi.pinned_empty if pinned else (lambda shape, dtype=np.float64: np.empty(shape, dtype))
        e, p = max(plan.nwin, 1), plan.npairs
        self.plan = plan
        self.mdccm, self.mad, self.vel, self.baz, self.sigma_tau = (mk((e,)) for _ in range(5))
        self.slowness = mk((e, 2))
        self.lag = mk((e, p), np.int32)
        self.cmax = mk((e, p)) if want_cmax else None
        self.weights = mk((e, p), np.uint8) if plan.alpha != 1.0 else None
        o = _capi.Outputs()
        for name in ("mdccm", "mad", "vel", "baz", "sigma_tau", "slowness", "lag", "cmax", "weights"):
            a = getattr(self, name)
            setattr(o, name, a.ctypes.data if a is not None else None)
        self.struct = o

    def band(self, b):
        """Per-band views in the reference's shapes: tau (P, W), element_weights (P, W), ..."""
        r = self.plan.band_rows(b)
        fs = self.plan.fs
        d = dict(vel=self.vel[r], baz=self.baz[r], mdccm=self.mdccm[r], mad=self.mad[r], sigma_tau=self.sigma_tau[r],
                 slowness=self.slowness[r].T, lag=self.lag[r].T, tau=(-self.lag[r].T.astype(np.float64)) / fs)
        if self.cmax is not None:
            d["cmax"] = self.cmax[r].T
        if self.weights is not None:
            d["element_weights"] = self.weights[r].T
        return d

    def nbytes(self):
        return sum(a.nbytes for a in (self.mdccm, self.mad, self.vel, self.baz, self.sigma_tau, self.slowness,
                                      self.lag, self.cmax, self.weights) if a is not None)


class DeviceArray:
    def __init__(self, eng, nbytes):
        self.eng, self.nbytes = eng, int(nbytes)
        p = C.c_void_p()
        check(eng.lib.nbls_malloc_device(eng.ctx, C.byref(p), max(self.nbytes, 1)))
        self.ptr = p.value

    def upload(self, arr):
        arr = np.ascontiguousarray(arr)
        check(self.eng.lib.nbls_memcpy_h2d(self.eng.ctx, self.ptr, arr.ctypes.data, arr.nbytes))
        return self

    def download(self, shape, dtype):
        out = np.empty(shape, dtype=dtype)
        check(self.eng.lib.nbls_memcpy_d2h(self.eng.ctx, out.ctypes.data, self.ptr, out.nbytes))
        return out

    def free(self):
        if self.ptr:
            self.eng.lib.nbls_free_device(self.eng.ctx, self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Engine:
    def __init__(self, device=0):
        self.lib = _capi.load()
        ctx = C.c_void_p()
        check(self.lib.nbls_create(C.byref(ctx), int(device)))
        self.ctx = ctx
        self.device = int(device)

    def close(self):
        if getattr(self, "ctx", None):
            self.lib.nbls_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- whole path -------------------------------------------------------------------------
    def run(self, plan: Plan, data, results: Results | None = None) -> Results:
        """Host buffers in, host buffers out (nbls_run_host): H2D + filter + xcorr + solve + D2H."""
        data = np.ascontiguousarray(data, dtype=np.float64)
        if data.shape != (plan.nchan, plan.npts):
            raise ValueError(f"data must have shape ({plan.nchan}, {plan.npts})")
        res = results if results is not None else Results(plan, pinned=False)
        check(self.lib.nbls_run_host(self.ctx, C.byref(plan.job), data.ctypes.data, plan.nchan, plan.npts,
                                     data.shape[1], C.byref(res.struct)))
        return res

    def resolve(self, plan: Plan, results: Results | None = None) -> Results:
        """Solve again on the lags this context kept from its last ``run`` with ``plan``'s solver description
        (nbls_resolve_host) — the second leg of an OLS + LTS sweep.  ``plan`` must describe the same record and window
        grid as the run (use ``Plan.with_alpha``)."""
        res = results if results is not None else Results(plan, pinned=False, want_cmax=False)
        check(self.lib.nbls_resolve_host(self.ctx, C.byref(plan.job.solve), C.byref(res.struct)))
        return res

    def run_resident(self, plan: Plan, x_dev: DeviceArray, ldx, out_struct):
        """Device buffers in and out (nbls_run_device); asynchronous — call synchronize()."""
        check(self.lib.nbls_run_device(self.ctx, C.byref(plan.job), x_dev.ptr, plan.nchan, plan.npts, int(ldx),
                                       C.byref(out_struct)))

    def device_outputs(self, plan: Plan):
        """Allocate a device-side output block; returns (Outputs struct, dict of DeviceArrays)."""
        e, p = max(plan.nwin, 1), plan.npairs
        sizes = dict(mdccm=8 * e, mad=8 * e, vel=8 * e, baz=8 * e, sigma_tau=8 * e, slowness=16 * e, lag=4 * e * p,
                     cmax=8 * e * p)
        if plan.alpha != 1.0:
            sizes["weights"] = e * p
        arrs = {k: DeviceArray(self, v) for k, v in sizes.items()}
        o = _capi.Outputs()
        for k, a in arrs.items():
            setattr(o, k, a.ptr)
        return o, arrs

    def synchronize(self):
        check(self.lib.nbls_synchronize(self.ctx))

    def timings(self):
        """Device ms of the last run: dict(h2d, filter, xcorr, solve, d2h, total)."""
        buf = (C.c_float * 6)()
        n = self.lib.nbls_last_timings(self.ctx, buf, 6)
        if n != 6:
            raise NBLSError(self.lib.nbls_last_error().decode())
        return dict(zip(("h2d", "filter", "xcorr", "solve", "d2h", "total"), (float(v) for v in buf)))

    def mark(self, slot):
        """Time stamp on the context's stream (nbls_mark)."""
        check(self.lib.nbls_mark(self.ctx, int(slot)))

    def elapsed(self, a, b):
        """Device ms between two marks; waits for the second."""
        v = C.c_float()
        check(self.lib.nbls_mark_elapsed(self.ctx, int(a), int(b), C.byref(v)))
        return float(v.value)

    def solve_resident(self, plan: Plan, nwin, lag_ptr, cmax_ptr, out_struct):
        """nbls_solve on device-resident lags / cmax with ``plan``'s solver description; asynchronous."""
        check(self.lib.nbls_solve(self.ctx, C.byref(plan.job.solve), int(nwin), lag_ptr, cmax_ptr, C.byref(out_struct)))

    def experiment_dft_pass(self, x, mode, reps=1, iters=3):
        """Tensor-pipe experiment (nbls_experiment_dft_pass): x (rows, 32) complex64 -> (y (rows, 32) complex64, ms)."""
        x = np.ascontiguousarray(x, dtype=np.complex64)
        rows = x.shape[0]
        xd = DeviceArray(self, x.nbytes).upload(x)
        yd = DeviceArray(self, x.nbytes)
        ms = C.c_float()
        try:
            check(self.lib.nbls_experiment_dft_pass(self.ctx, int(mode), xd.ptr, rows, int(reps), int(iters), yd.ptr,
                                                    C.byref(ms)))
            self.synchronize()
            y = yd.download(x.shape, np.complex64)
        finally:
            xd.free(); yd.free()
        return y, float(ms.value)

    def launch_count(self):
        return int(self.lib.nbls_launch_count(self.ctx))

    def fp64_peak_tflops(self, iters=20000):
        v = C.c_double()
        check(self.lib.nbls_measure_fp64_peak(self.ctx, int(iters), C.byref(v)))
        return v.value

    def fp32_peak_tflops(self, iters=20000):
        v = C.c_double()
        check(self.lib.nbls_measure_fp32_peak(self.ctx, int(iters), C.byref(v)))
        return v.value

    def xcorr_stats(self, reset=False):
        """Diagnostics of the FP32-screen cross-correlation (nbls_xcorr_stats): candidates verified exactly, items the
        time-domain kernel settled, largest |FP32 screen maximum - exact maximum| seen."""
        nv, nt, err = C.c_int64(), C.c_int64(), C.c_double()
        check(self.lib.nbls_xcorr_stats(self.ctx, C.byref(nv), C.byref(nt), C.byref(err), int(bool(reset))))
        nw, nr = C.c_int64(), C.c_int64()
        check(self.lib.nbls_xcorr_two_stage_stats(self.ctx, C.byref(nw), C.byref(nr), int(bool(reset))))
        npr = C.c_int64()
        check(self.lib.nbls_xcorr_two_stage_pairs(self.ctx, C.byref(npr), int(bool(reset))))
        return dict(verified=nv.value, time_domain=nt.value, max_screen_err=err.value, two_stage_windows=nw.value,
                    two_stage_redone=nr.value, two_stage_redone_pairs=npr.value)

    def xcorr_audit(self, reset=False):
        """(largest |FP32 screen - FP64 FFT| over every searched lag, lags compared) while OPT_SCREEN_AUDIT was on."""
        err, n = C.c_double(), C.c_int64()
        check(self.lib.nbls_xcorr_audit(self.ctx, C.byref(err), C.byref(n), int(bool(reset))))
        return err.value, n.value

    def set_option(self, key, value):
        """Per-context switch (nbls_set_option; keys are _capi.OPT_*)."""
        check(self.lib.nbls_set_option(self.ctx, int(key), int(value)))

    def get_option(self, key):
        v = C.c_int64()
        check(self.lib.nbls_get_option(self.ctx, int(key), C.byref(v)))
        return v.value

    def profile(self, reset=False):
        """Per-kernel-group device ms / launches / executed algorithmic work gathered while OPT_PROFILE was on."""
        out = {}
        for g in range(self.lib.nbls_kernel_group_count()):
            ms, n, wk = C.c_double(), C.c_int64(), C.c_double()
            check(self.lib.nbls_profile(self.ctx, g, C.byref(ms), C.byref(n), C.byref(wk), int(bool(reset))))
            out[self.lib.nbls_kernel_group_name(g).decode()] = dict(ms=ms.value, launches=n.value, work=wk.value)
        return out

    # ---- stage level ------------------------------------------------------------------------
    def filter_bands(self, data, fs, sos_list, zerophase=True, taper=0.01, chunk_len=0, exact_scan=False,
                     global_t0=0, global_npts=None):
        """(C, T) -> (B, C, T) filtered + tapered, one launch sequence for all bands (nbls_filter)."""
        data = np.ascontiguousarray(data, dtype=np.float64)
        c, t = data.shape
        sos = np.ascontiguousarray(np.stack([np.asarray(s, dtype=np.float64) for s in sos_list]))
        b, s = sos.shape[:2]
        ld = (t + 15) // 16 * 16
        g = global_npts if global_npts is not None else t
        wlen, up, down = taper_ramps(g, taper) if taper else (0, np.zeros(0), np.zeros(0))
        d = _capi.FilterDesc()
        d.nbands, d.nsections, d.sos = b, s, _capi.dptr(sos)
        d.zerophase, d.exact_scan, d.chunk_len = int(bool(zerophase)), int(bool(exact_scan)), int(chunk_len)
        d.taper_len, d.taper_up, d.taper_down = wlen, _capi.dptr(up), _capi.dptr(down)
        d.global_t0, d.global_npts = int(global_t0), int(g)
        xpad = np.zeros((c, ld))
        xpad[:, :t] = data
        xd = DeviceArray(self, xpad.nbytes).upload(xpad)
        yd = DeviceArray(self, 8 * b * c * ld)
        try:
            check(self.lib.nbls_filter(self.ctx, C.byref(d), xd.ptr, c, t, ld, yd.ptr, ld, c * ld))
            self.synchronize()
            out = yd.download((b, c, ld), np.float64)[:, :, :t]
        finally:
            xd.free()
            yd.free()
        return np.ascontiguousarray(out)

    def xcorr(self, filt, win_len, win_hop, win_count, win_first=None, max_lag=-1, method=_capi.XCORR_AUTO, want_cmax=True):
        """filt (B, C, T) -> lag (E, P) int32, cmax (E, P)  (nbls_xcorr).  ``want_cmax=False`` passes no cmax buffer (the
        FP32-screen path then evaluates only undecided pair-windows in FP64, NBLS_OPT_LAZY_VERIFY) and returns cmax = None."""
        filt = np.asarray(filt, dtype=np.float64)
        b, c, t = filt.shape
        ld = (t + 15) // 16 * 16
        ypad = np.zeros((b, c, ld))
        ypad[:, :, :t] = filt
        wl = np.ascontiguousarray(win_len, dtype=np.int32)
        wh = np.ascontiguousarray(win_hop, dtype=np.int32)
        wc = np.ascontiguousarray(win_count, dtype=np.int32)
        wf = np.ascontiguousarray(win_first if win_first is not None else np.zeros(b), dtype=np.int64)
        w = _capi.WindowDesc()
        w.nbands, w.max_lag, w.method = b, int(max_lag), int(method)
        w.win_len, w.win_hop = _capi.dptr(wl, _capi.c_int32_p), _capi.dptr(wh, _capi.c_int32_p)
        w.win_first, w.win_count = _capi.dptr(wf, _capi.c_int64_p), _capi.dptr(wc, _capi.c_int32_p)
        e, p = int(wc.sum()), c * (c - 1) // 2
        yd = DeviceArray(self, ypad.nbytes).upload(ypad)
        ld_ = DeviceArray(self, 4 * max(e, 1) * p)
        cd = DeviceArray(self, 8 * max(e, 1) * p)
        try:
            check(self.lib.nbls_xcorr(self.ctx, C.byref(w), yd.ptr, c, ld, c * ld, ld_.ptr, cd.ptr if want_cmax else None))
            self.synchronize()
            lag = ld_.download((e, p), np.int32)
            cmax = cd.download((e, p), np.float64) if want_cmax else None
        finally:
            yd.free(); ld_.free(); cd.free()
        return lag, cmax

    def solve(self, lag, cmax, rij, fs, alpha=1.0):
        """lag/cmax (E, P) -> dict of per-window estimates (nbls_solve)."""
        lag = np.ascontiguousarray(lag, dtype=np.int32)
        cmax = np.ascontiguousarray(cmax, dtype=np.float64)
        e, p = lag.shape
        co = rij if isinstance(rij, CoArray) else CoArray(np.asarray(rij, dtype=np.float64))
        plan = Plan(co.nchans, 1, fs, np.zeros((1, 1, 6)), True, [1], [1], [0], co, alpha=alpha, taper=0)
        o, arrs = self.device_outputs_for(e, p, alpha != 1.0)
        arrs["lag"].upload(lag)
        arrs["cmax"].upload(cmax)
        try:
            check(self.lib.nbls_solve(self.ctx, C.byref(plan.job.solve), e, arrs["lag"].ptr, arrs["cmax"].ptr, C.byref(o)))
            self.synchronize()
            out = {k: arrs[k].download((e,), np.float64) for k in ("mdccm", "mad", "vel", "baz", "sigma_tau")}
            out["slowness"] = arrs["slowness"].download((e, 2), np.float64)
            if alpha != 1.0:
                out["weights"] = arrs["weights"].download((e, p), np.uint8)
        finally:
            for a in arrs.values():
                a.free()
        return out

    def device_outputs_for(self, e, p, lts):
        e = max(int(e), 1)
        sizes = dict(mdccm=8 * e, mad=8 * e, vel=8 * e, baz=8 * e, sigma_tau=8 * e, slowness=16 * e, lag=4 * e * p,
                     cmax=8 * e * p)
        if lts:
            sizes["weights"] = e * p
        arrs = {k: DeviceArray(self, v) for k, v in sizes.items()}
        o = _capi.Outputs()
        for k, a in arrs.items():
            setattr(o, k, a.ptr)
        return o, arrs


_DEFAULT = {}


def default_engine(device=None) -> Engine:
    """Process-wide engine per device (LOCAL_RANK selects the device under torchrun)."""
    import os
    if device is None:
        device = int(os.environ.get("LOCAL_RANK", "0"))
    if device not in _DEFAULT:
        _DEFAULT[device] = Engine(device)
    return _DEFAULT[device]
