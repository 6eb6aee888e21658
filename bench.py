#!/usr/bin/env python
"""bench.py — NBLS band-window estimates per second on B200 (BASELINE.json metric).

  python bench.py [--gpus N --steps K --warmup W] [--impl reference] [--workload c1|c2|c3|c4|c5] [--hours H]
                  [--scaling weak|strong] [--max-lag L] [--method auto|fft32|fft64|direct]

One "step" = one pass of the whole hot path (per-band zero-phase band-pass -> windowed pairwise cross-correlation ->
slowness solve) over the synthetic record(s) of the workload.  Default c3: 25 elements, 24 h at 100 sps, 20 log bands
0.5-5 Hz, OLS, adaptive 60->30 s windows, 50 % overlap.
  --scaling weak   (default) N > 1: the record is N x the workload's length, sharded by time segment with filter halos.
  --scaling strong N > 1: ONE record of the workload's length (c4: 30 days) is split over the N ranks.
  c5               a batch of independent arrays (default 64), dealt to the ranks by array (batch.deal_arrays), each
                   array an OLS + LTS sweep on one filter + cross-correlation pass.
Per-window estimates are gathered with ONE NCCL all_gather_into_tensor of a packed block (sharding.PackedGather).

JSON line:
 `value`        estimates / device time, inputs resident in HBM (CUDA events on libnbls' stream, max over ranks).
 `e2e`          the same through the host-buffer C-ABI call nbls_run_host (pinned H2D + path + D2H inside the timed region),
                by default with two records in flight on two contexts, which is why it can exceed `value`: the tail of one
                record overlaps the head of the next (`ms_record_latency_device` is the time ONE record spends in flight).
 `roofline`     the dominant kernel (largest device time of a profiled step; normally xf_pair_kernel): EXECUTED
                algorithmic flop of its launches / its CUDA-event time, against the FP32 FMA peak measured in-process
                (the kernel is issue-bound on the FP32 pipe; neither HBM nor the tensor pipe binds it — DESIGN.md §4).
 `kernel_groups` every kernel group of that profiled step with its own time, executed work and fraction of ITS bound
                (filter groups: HBM, MEASURED_PEAKS.json; FFT groups: FP32 FMA peak; exact verification: FP64 FMA peak).
 `roofline_filter` the HBM-bound filter stage as a whole (algorithmic bytes / stage time of the timed steps).
 `parity`       the GPU path against the CPU oracle AND the reference's own driver on the same bounded slice the CPU
                baseline is timed on: integer-lag mismatches, NaN pattern, max deviations of the estimates.
 `cpu_baseline` / --impl reference: the UNMODIFIED reference's narrow_band_least_squares_parallel (joblib over bands;
                staged to oracle/_ref by build()) through the import shims of oracle/refstubs, on the box's host cores,
                bounded slice, cold (first call: pool start, numba JIT) and warm; falls back to the oracle port only if
                the staged reference is absent.
"""
from __future__ import annotations

import argparse
import contextlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

FALLBACK_HBM_GBS = 6650.0


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c3")
    ap.add_argument("--hours", type=float, default=None, help="record length (default: the workload's)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--narrays", type=int, default=None, help="c5: arrays in the batch (default 64)")
    ap.add_argument("--alpha", type=float, default=None, help="override the workload's ALPHA (1.0 = OLS, [0.5, 1) = FAST-LTS)")
    ap.add_argument("--method", default="auto", choices=["auto", "direct", "fft", "fft64", "fft32"])
    ap.add_argument("--max-lag", type=int, default=-1, help="restrict |lag| (default: all 2N-1 lags = reference)")
    ap.add_argument("--single-stage", action="store_true", help="switch the two-stage lag search off")
    ap.add_argument("--noise-only", action="store_true", help="incoherent-noise record (no plane wave)")
    ap.add_argument("--cpu-sample-min", type=float, default=3.0, help="minutes of record for the CPU legs")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline and parity legs")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-profile", action="store_true", help="skip the profiled step (kernel_groups)")
    ap.add_argument("--opt", action="append", default=[], metavar="KEY=VALUE",
                    help="libnbls context option by number (include/nbls.h NBLS_OPT_*), e.g. --opt 7=4")
    ap.add_argument("--e2e-in-flight", type=int, default=3,
                    help="records in flight in the e2e leg (own context and workspace each, capped by what fits in HBM; "
                         "1 = strictly one after the other)")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi sampling during the timed region (B200_PROFILING.md clocks line)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc, self.t_mark = index, [], None, None

    def mark(self):
        """Start of the timed region: only samples read after this count (nvidia-smi is started before the warm-up
        steps so that its start-up does not stall the first timed step)."""
        self.t_mark = time.perf_counter()

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.index)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [c.strip() for c in line.split(",")]))

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        rows = [r for t, r in self.rows if self.t_mark is None or t >= self.t_mark]
        if not rows:                      # timed region shorter than one sampling period: take the last sample
            rows = [r for _, r in self.rows[-1:]]
        for r in rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2])); pw.append(float(r[3]))
            except Exception:
                continue
            for name, col in (("hw_slowdown", 5), ("hw_thermal_slowdown", 6), ("sw_thermal_slowdown", 7),
                              ("sw_power_cap", 8)):
                if len(r) > col and r[col].lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f), "MEASURED_PEAKS.json hbm_gbs (burst copy figure)"
    return {"hbm_gbs": FALLBACK_HBM_GBS}, "fallback of B200_PROFILING.md (MEASURED_PEAKS.json absent)"


def ncu_traffic(key):
    """DRAM bytes per step of each kernel group from the committed ncu pass (scripts/ncu_traffic.sh ->
    profiles/ncu_traffic.json), or {} when no capture exists for this workload key."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    try:
        with open(p) as f:
            return json.load(f).get(key, {})
    except Exception:
        return {}


def workload(args):
    """Host-side description of the job: config dict, per-GPU npts."""
    from narrow_band_least_squares_b200.configs import config_job
    j = config_job(args.workload)
    j["hours_total"] = j["hours"] if args.hours is None else args.hours
    if args.alpha is not None:
        j["alpha"] = args.alpha
    j["narrays"] = (args.narrays or 64) if args.workload == "c5" else 1
    return j


def make_record(j, npts, seed_offset=0, noise_only=False):
    from narrow_band_least_squares_b200.synth import CONFIGS, plane_wave_record
    c = CONFIGS[j["name"]]
    kw = dict(flip_channel=c.get("flip_channel"), shift_channel=c.get("shift_channel"))
    if noise_only:
        kw.update(sigma=30.0, noise_every_min=0)        # the plane wave is 30 dB below the incoherent noise
    return plane_wave_record(c["seed"] + seed_offset, c["nchans"], npts, c["fs"], c["aperture_km"], **kw)


@contextlib.contextmanager
def quiet_stdout():
    """The reference prints per band and per ltsva call (also from joblib workers): send fd 1 to /dev/null meanwhile so
    that the JSON line stays the only thing on stdout."""
    sys.stdout.flush()
    saved = os.dup(1)
    devnull = os.open(os.devnull, os.O_WRONLY)
    os.dup2(devnull, 1)
    try:
        yield
    finally:
        sys.stdout.flush()
        os.dup2(saved, 1)
        os.close(saved)
        os.close(devnull)


def cpu_port_run(j, rec, n_jobs):
    """The oracle port (joblib over bands) on ``rec``: outputs with integer lags, for the parity block."""
    from oracle import nbls_oracle as orc
    tvec = np.arange(rec.npts) / j["fs"] / 86400.0
    t0 = time.perf_counter()
    out = orc.nbls_oracle(rec.data, j["fs"], tvec, rec.rij, j["freqlist"], j["winlen_list"], j["winover"], j["alpha"],
                          ftype=j["filter_type"], order=j["filter_order"], ripple=j["filter_ripple"], n_jobs=n_jobs)
    return out, time.perf_counter() - t0


def cpu_reference_run(j, rec):
    """The reference's own narrow_band_least_squares_parallel (narrow_band_least_squares.py:223-323, joblib n_jobs=-1
    at :285) on ``rec`` through the import shims.  Returns (9-tuple, seconds) or None if no reference tree exists."""
    from oracle import refrun
    if refrun.reference_dir() is None:
        return None
    ref = refrun.import_reference()
    with quiet_stdout():
        t0 = time.perf_counter()
        out = refrun.run_reference_driver(ref, rec, j["freqlist"], j["winlen_list"], j["winover"], j["alpha"],
                                          j["nbands"], ftype=j["filter_type"], order=j["filter_order"],
                                          ripple=j["filter_ripple"], parallel=True)
        dt = time.perf_counter() - t0
    return out, dt


def cpu_legs(j, minutes, ncores, want_port=True):
    """CPU baseline on a bounded slice: the reference driver cold + warm (kind "reference"), else the port; and the
    port's outputs (with lags) for the parity block.  Returns (cpu_baseline dict, rec, ref_out, port_out)."""
    npts = int(round(minutes * 60.0 * j["fs"]))
    rec = make_record(j, npts)
    ref_out = port_out = None
    r = cpu_reference_run(j, rec)
    if r is not None:
        ref_out, cold = r
        ref_out, warm = cpu_reference_run(j, rec)
        est = int(sum(ref_out[6]))
        base = {"value": est / warm, "unit": "estimates/s", "cores": ncores, "kind": "reference",
                "cold_value": est / cold,
                "sample": (f"{minutes:g}-minute slice of the workload, all {j['nbands']} bands, {est} estimates: the "
                           f"reference's narrow_band_least_squares_parallel (joblib n_jobs=-1) took {cold:.1f} s cold "
                           f"(pool start + numba JIT in every worker) and {warm:.1f} s warm; `value` is the warm rate")}
    if want_port or r is None:
        port_out, dt = cpu_port_run(j, rec, ncores)
        if r is None:
            est = int(sum(port_out["num_compute"]))
            base = {"value": est / dt, "unit": "estimates/s", "cores": ncores, "kind": "port",
                    "sample": f"{minutes:g}-minute slice, all {j['nbands']} bands, {est} estimates in {dt:.1f} s "
                              "(oracle port: the staged reference oracle/_ref is absent)"}
    return base, rec, ref_out, port_out


def _baz_diff(a, b):
    return np.abs(((a - b + 180.0) % 360.0) - 180.0)


def parity_block(eng, j, rec, ref_out, port_out, args, method):
    """GPU path vs the CPU oracle (integer lags, flags, estimates) and vs the reference driver's arrays on ``rec``."""
    from narrow_band_least_squares_b200.configs import config_plan
    from narrow_band_least_squares_b200.engine import Results
    from oracle import nbls_oracle as orc
    plan = config_plan(j["name"], rec.npts, rec.rij, method=method, max_lag=args.max_lag)
    res = eng.run(plan, rec.data, Results(plan, pinned=False, want_cmax=True))
    P, nb = plan.npairs, plan.nbands
    out = {"sample": f"{rec.npts / j['fs'] / 60.0:g}-minute slice of the workload (the cpu_baseline record), all {nb} bands",
           "estimates": int(plan.nwin), "pair_windows": int(plan.nwin) * P}
    # the timed steps ask for no per-pair cmax (lazy verification, NBLS_OPT_LAZY_VERIFY); the comparison below runs WITH
    # cmax (every pair-window evaluated in FP64).  Both paths must agree: lags and estimates bit for bit, MdCCM to the last bit
    fast = eng.run(plan, rec.data, Results(plan, pinned=False, want_cmax=False))
    with np.errstate(invalid="ignore"):
        out["timed_path_vs_cmax_path"] = {
            "lag_mismatches": int((fast.lag != res.lag).sum()),
            "estimates_bit_identical": bool(all(np.array_equal(getattr(fast, k), getattr(res, k), equal_nan=True)
                                                for k in ("mad", "vel", "baz", "sigma_tau"))),
            "max_abs_dmdccm": float(np.nanmax(np.abs(fast.mdccm - res.mdccm), initial=0.0)),
            "lts_flags_equal": (bool(np.array_equal(fast.weights, res.weights)) if res.weights is not None else None)}
    if port_out is not None and args.max_lag < 0:
        mism, mism_filt, nan_ok, dvel, dbaz, dmd, dsig, dcmax, flag_mism = 0, None, True, 0.0, 0.0, 0.0, 0.0, 0.0, 0
        for b in range(nb):
            g, e = res.band(b), port_out["bands"][b]
            bad = g["lag"] != e["lag"]
            nbad = int(bad.sum())
            mism += nbad
            if nbad:
                # classify: the same cross-correlation on the ORACLE-filtered samples - what is left then is not a
                # near-tie flipped by the 1e-12 difference between the GPU's chunked filter scan and scipy's sosfilt
                filt, _ = orc.filter_record(rec.data, j["fs"], j["filter_type"], j["freqlist"][b], j["freqlist"][b + 1],
                                            j["filter_order"], j["filter_ripple"])
                lag2, _ = eng.xcorr(filt[None], [int(plan.win_len[b])], [int(plan.win_hop[b])], [int(plan.win_count[b])],
                                    method=method)
                mism_filt = (mism_filt or 0) + int((lag2.T != e["lag"]).sum())
            for k in ("vel", "baz", "mdccm", "sigma_tau"):
                nan_ok = nan_ok and bool(np.array_equal(np.isnan(g[k]), np.isnan(e[k])))
            with np.errstate(invalid="ignore", divide="ignore"):
                dvel = max(dvel, float(np.nanmax(np.abs(g["vel"] - e["vel"]) / np.abs(e["vel"]), initial=0.0)))
                dbaz = max(dbaz, float(np.nanmax(_baz_diff(g["baz"], e["baz"]), initial=0.0)))
                dmd = max(dmd, float(np.nanmax(np.abs(g["mdccm"] - e["mdccm"]), initial=0.0)))
                dsig = max(dsig, float(np.nanmax(np.abs(g["sigma_tau"] - e["sigma_tau"]) /
                                                 np.maximum(np.abs(e["sigma_tau"]), 1e-300), initial=0.0)))
                dcmax = max(dcmax, float(np.nanmax(np.abs(g["cmax"] - e["cmax"]), initial=0.0)))
            if j["alpha"] != 1.0:
                flag_mism += int(np.any(g["element_weights"] != e["element_weights"], axis=0).sum())
        out["vs_oracle"] = {"lag_mismatches": mism, "lag_mismatches_on_oracle_filtered_samples": mism_filt,
                            "nan_pattern_equal": nan_ok, "max_rel_dvel": dvel, "max_dbaz_deg": dbaz,
                            "max_abs_dmdccm": dmd, "max_rel_dsigma_tau": dsig, "max_abs_dcmax": dcmax}
        if j["alpha"] != 1.0:
            out["vs_oracle"]["lts_flag_window_mismatches"] = flag_mism
    if ref_out is not None and args.max_lag < 0:
        vel, baz, md, tt, _, sig, ncl = ref_out[:7]
        ok, dvel, dbaz, dmd, dsig = True, 0.0, 0.0, 0.0, 0.0
        for b in range(nb):
            w = int(ncl[b])
            g = res.band(b)
            ok = ok and w == int(plan.win_count[b]) and bool(np.array_equal(np.isnan(g["vel"]), np.isnan(vel[b, :w])))
            if w != int(plan.win_count[b]):
                continue
            with np.errstate(invalid="ignore", divide="ignore"):
                dvel = max(dvel, float(np.nanmax(np.abs(g["vel"] - vel[b, :w]) / np.abs(vel[b, :w]), initial=0.0)))
                dbaz = max(dbaz, float(np.nanmax(_baz_diff(g["baz"], baz[b, :w]), initial=0.0)))
                dmd = max(dmd, float(np.nanmax(np.abs(g["mdccm"] - md[b, :w]), initial=0.0)))
                if j["alpha"] == 1.0:      # the reference leaves sig_tau unset for ALPHA < 1
                    dsig = max(dsig, float(np.nanmax(np.abs(g["sigma_tau"] - sig[b, :w]) /
                                                     np.maximum(np.abs(sig[b, :w]), 1e-300), initial=0.0)))
        out["vs_reference_driver"] = {"window_counts_and_nan_pattern_equal": ok, "max_rel_dvel": dvel,
                                      "max_dbaz_deg": dbaz, "max_abs_dmdccm": dmd, "max_rel_dsigma_tau": dsig}
    out["tolerances"] = "lags/flags exact; vel, sigma_tau 1e-6 rel; baz 1e-3 deg; mdccm 1e-9 (north_star)"
    return out


class _DevPtr:
    """A device buffer owned by a torch tensor, in the shape Engine.run_resident expects (``.ptr``)."""

    def __init__(self, tensor):
        self.tensor, self.ptr = tensor, tensor.data_ptr()


def dist_setup():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    return world, rank, local_rank, dist


def bind_to_gpu_numa_node(local_rank):
    """Multi-GPU e2e: run this rank's host threads - and therefore place its pinned buffers (first touch) - on the NUMA
    node its GPU hangs off, so that eight ranks do not pull their H2D copies through one memory controller.  Returns the
    node (or None when sysfs does not say)."""
    try:
        import torch
        pr = torch.cuda.get_device_properties(local_rank)
        dev = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        with open(f"/sys/bus/pci/devices/{dev}/numa_node") as f:
            node = int(f.read().strip())
        if node < 0:
            return None
        with open(f"/sys/devices/system/node/node{node}/cpulist") as f:
            cpus = set()
            for part in f.read().strip().split(","):
                a, _, b = part.partition("-")
                cpus.update(range(int(a), int(b or a) + 1))
        os.sched_setaffinity(0, cpus & os.sched_getaffinity(0) or cpus)
        return node
    except Exception:
        return None


def xcorr_method(args):
    from narrow_band_least_squares_b200 import _capi
    return {"auto": _capi.XCORR_AUTO, "direct": _capi.XCORR_DIRECT, "fft": _capi.XCORR_FFT, "fft64": _capi.XCORR_FFT,
            "fft32": _capi.XCORR_FFT32}[args.method]


def segment_record(j, args, seg_lo, seg_hi, ld, local_rank, world):
    """This rank's samples [seg_lo, seg_hi) of the (global) record: (device buffer [C, ld], pinned host copy, rij).
    One GPU and a record numpy can draw in seconds: the round-1 generator (same data as BENCH_r01); otherwise the
    block-wise generator on the GPU (synth.blocked_record_cuda), identical on every rank for a given sample index."""
    from narrow_band_least_squares_b200 import _capi
    from narrow_band_least_squares_b200.synth import CONFIGS, blocked_record_cuda
    c = CONFIGS[j["name"]]
    npts = seg_hi - seg_lo
    x_host = _capi.pinned_empty((c["nchans"], npts))
    import torch
    dev = torch.device("cuda", local_rank)
    if world == 1 and npts * c["nchans"] <= 300e6:
        rec = make_record(j, npts, noise_only=args.noise_only)
        x_host[:] = rec.data
        x_dev = torch.zeros((c["nchans"], ld), dtype=torch.float64, device=dev)
        x_dev[:, :npts] = torch.from_numpy(x_host)
        rij = rec.rij
    else:
        kw = dict(flip_channel=c.get("flip_channel"), shift_channel=c.get("shift_channel"))
        if args.noise_only:
            kw.update(sigma=30.0, noise_every_min=0)
        x_dev, rij, _, _ = blocked_record_cuda(c["seed"], c["nchans"], seg_lo, seg_hi, c["fs"], c["aperture_km"], dev,
                                               ld=ld, **kw)
        torch.from_numpy(x_host).copy_(x_dev[:, :npts])
    torch.cuda.synchronize()
    torch.cuda.empty_cache()
    return _DevPtr(x_dev), x_host, rij


def multi_gpu_check(j, eng, rank, world, dist, sos, method, args):
    """The NCCL path against an expected value: a 2-hour block-wise record, every rank runs its time shard and the
    estimates, integer lags (and LTS flags) are gathered; rank 0 also runs the record unsharded on its own GPU."""
    from narrow_band_least_squares_b200.engine import Plan, Results, window_grid
    from narrow_band_least_squares_b200.sharding import PackedGather, Shard, all_shards, filter_halo
    from narrow_band_least_squares_b200.synth import CONFIGS, blocked_record
    import torch
    c = CONFIGS[j["name"]]
    fs, nb = j["fs"], j["nbands"]
    g_npts = int(2 * 3600 * fs)
    grids = [window_grid(g_npts, fs, j["winlen_list"][b], j["winover"]) for b in range(nb)]
    n = [g[0] for g in grids]; hop = [g[1] for g in grids]; cnt = [g[2] for g in grids]
    halo = filter_halo(sos)
    shards = all_shards(g_npts, n, hop, cnt, world, halo)
    sh = shards[rank]
    kw = dict(flip_channel=c.get("flip_channel"), shift_channel=c.get("shift_channel"))
    rec = blocked_record(c["seed"] + 99, c["nchans"], sh.seg_lo, sh.seg_hi, fs, c["aperture_km"], **kw)
    plan = sh.plan(c["nchans"], fs, sos, j["filter_type"] == "butter", n, hop, rec.rij, alpha=j["alpha"], method=method)
    res = eng.run(plan, rec.data, Results(plan, pinned=False, want_cmax=False))
    lts = j["alpha"] != 1.0
    pg = PackedGather([s.win_count for s in shards], rank, npairs=plan.npairs, want_lag=True, want_weights=lts)
    e = plan.nwin
    for i, k in enumerate(("mdccm", "mad", "vel", "baz", "sigma_tau")):
        pg.local[i, :e] = torch.from_numpy(getattr(res, k)[:e]).cuda()
    pg.local_lag[:e] = torch.from_numpy(res.lag[:e]).cuda()
    if lts:
        pg.local_w[:e] = torch.from_numpy(res.weights[:e]).cuda()
    got = {k: v.cpu().numpy() for k, v in pg.gather().items()}
    if rank != 0:
        return None
    whole = blocked_record(c["seed"] + 99, c["nchans"], 0, g_npts, fs, c["aperture_km"], **kw)
    wplan = Plan(c["nchans"], g_npts, fs, sos, j["filter_type"] == "butter", n, hop, cnt, whole.rij, alpha=j["alpha"],
                 method=method)
    ref = eng.run(wplan, whole.data, Results(wplan, pinned=False, want_cmax=False))
    E = wplan.nwin
    with np.errstate(invalid="ignore", divide="ignore"):
        out = {"record": f"2 h block-wise {j['name']} record, {world} time shards with {halo}-sample filter halos over NCCL "
                         "vs the same record unsharded on rank 0",
               "estimates": int(E), "pair_windows": int(E) * plan.npairs,
               "lag_mismatches": int((got["lag"] != ref.lag[:E]).sum()),
               "nan_pattern_equal": bool(np.array_equal(np.isnan(got["vel"]), np.isnan(ref.vel[:E]))),
               "max_rel_dvel": float(np.nanmax(np.abs(got["vel"] - ref.vel[:E]) / np.abs(ref.vel[:E]), initial=0.0)),
               "max_dbaz_deg": float(np.nanmax(_baz_diff(got["baz"], ref.baz[:E]), initial=0.0))}
    if lts:
        out["lts_flag_window_mismatches"] = int(np.any(got["weights"] != ref.weights[:E], axis=1).sum())
    return out


def run_single(args):
    """c1-c4: one array, one (long) record; N > 1 shards it by time segment."""
    from narrow_band_least_squares_b200 import _capi
    from narrow_band_least_squares_b200.engine import Engine, Results, window_grid
    from narrow_band_least_squares_b200.helpers import band_edges, design_sos
    from narrow_band_least_squares_b200.sharding import PackedGather, all_shards, filter_halo

    world, rank, local_rank, dist = dist_setup()
    numa = bind_to_gpu_numa_node(local_rank) if world > 1 else None
    j = workload(args)
    method = xcorr_method(args)
    fs, nchan, nb = j["fs"], j["nchans"], j["nbands"]
    npts_total = int(round(j["hours_total"] * 3600.0 * fs))
    g_npts = npts_total * world if args.scaling == "weak" else npts_total
    sos = np.stack([design_sos(j["filter_type"], *band_edges(j["freqlist"], b, "log"), fs, j["filter_order"],
                               j["filter_ripple"]) for b in range(nb)])
    grids = [window_grid(g_npts, fs, j["winlen_list"][b], j["winover"]) for b in range(nb)]
    win_len = [g[0] for g in grids]; win_hop = [g[1] for g in grids]; win_cnt = [g[2] for g in grids]
    halo = filter_halo(sos) if world > 1 else 0
    shards = all_shards(g_npts, win_len, win_hop, win_cnt, world, halo)
    shard = shards[rank]
    ld = (shard.npts + 15) // 16 * 16
    eng = Engine(local_rank)
    if args.single_stage:
        eng.set_option(_capi.OPT_TWO_STAGE, 0)
    for kv in args.opt:
        eng.set_option(int(kv.split("=")[0]), int(kv.split("=")[1]))
    x_dev, x_host, rij = segment_record(j, args, shard.seg_lo, shard.seg_hi, ld, local_rank, world)
    plan = shard.plan(nchan, fs, sos, j["filter_type"] == "butter", win_len, win_hop, rij, alpha=j["alpha"],
                      max_lag=args.max_lag, method=method)
    e_local, e_total = plan.nwin, int(sum(win_cnt))
    lts = plan.alpha != 1.0
    res = Results(plan, pinned=True, want_cmax=False)
    pg = None
    if world > 1:
        import torch
        pg = PackedGather([s.win_count for s in shards], rank, npairs=plan.npairs, want_weights=lts)
        lag_t = torch.empty((max(e_local, 1), plan.npairs), dtype=torch.int32, device="cuda")
        o_dev = pg.outputs(lag_ptr=lag_t.data_ptr())
        keep = lag_t
    else:
        o_dev, keep = eng.device_outputs(plan)
        o_dev.cmax = None

    def barrier():
        eng.synchronize()
        if dist is not None:
            import torch
            torch.cuda.synchronize()
            dist.barrier()

    def step_resident():
        eng.run_resident(plan, x_dev, ld, o_dev)
        eng.synchronize()
        t = eng.timings()
        ms_gather = 0.0
        if pg is not None:
            import torch
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            pg.gather()
            e1.record()
            torch.cuda.synchronize()
            ms_gather = e0.elapsed_time(e1)
        return t, ms_gather

    fp64_peak = eng.fp64_peak_tflops(20000)
    fp32_peak = eng.fp32_peak_tflops(20000)
    sampler = ClockSampler(local_rank).start()
    for _ in range(args.warmup):
        step_resident()
    barrier()
    launches0 = eng.launch_count()
    eng.xcorr_stats(reset=True)
    sampler.mark()
    wall0 = time.perf_counter()
    dev_ms, gather_ms, stage = 0.0, 0.0, {"filter": 0.0, "xcorr": 0.0, "solve": 0.0}
    for _ in range(args.steps):
        t, mg = step_resident()
        dev_ms += t["total"] + mg
        gather_ms += mg
        for k in stage:
            stage[k] += t[k]
    barrier()
    wall_ms = (time.perf_counter() - wall0) * 1e3
    clocks = sampler.stop()
    launches = eng.launch_count() - launches0
    xstats = eng.xcorr_stats()
    xstats["pair_windows"] = int(e_local * plan.npairs * args.steps)
    if dist is not None:
        import torch
        tt = torch.tensor([dev_ms, wall_ms, gather_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dev_ms, wall_ms, gather_ms = float(tt[0]), float(tt[1]), float(tt[2])
    ms_per_step = dev_ms / args.steps
    value = e_total / (ms_per_step * 1e-3)

    # one more step with the per-kernel-group profile on (outside the timed region: it synchronises after every group)
    prof = None
    if not args.no_profile:
        eng.set_option(_capi.OPT_PROFILE, 1)
        eng.profile(reset=True)
        eng.run_resident(plan, x_dev, ld, o_dev)
        eng.synchronize()
        prof = eng.profile(reset=True)
        eng.set_option(_capi.OPT_PROFILE, 0)

    e2e = None
    if not args.no_e2e:
        # every context in flight holds its own workspace (filtered bands of all B bands + record + ~8 GB of cross-
        # correlation buffers): as many as fit in 80 % of the HBM, at most --e2e-in-flight
        import torch
        per_ctx = 8.0 * nchan * ld * (nb + 1) + 8e9
        fit = int(0.8 * torch.cuda.get_device_properties(local_rank).total_memory / per_ctx)
        nfl = max(1, min(args.e2e_in_flight, args.steps, fit))
        engs = [eng] + [Engine(local_rank) for _ in range(nfl - 1)]
        for e_ in engs[1:]:
            if args.single_stage:
                e_.set_option(_capi.OPT_TWO_STAGE, 0)
            for kv in args.opt:
                e_.set_option(int(kv.split("=")[0]), int(kv.split("=")[1]))
        ress = [res] + [Results(plan, pinned=True, want_cmax=False) for _ in range(nfl - 1)]
        for e_, r_ in zip(engs, ress):
            e_.run(plan, x_host, r_)                          # one warm-up each (buffers grow once)
        for r_ in ress[1:]:                                   # every context computes the same bits
            assert np.array_equal(r_.lag, res.lag) and np.array_equal(r_.vel, res.vel, equal_nan=True)
        barrier()
        done = [threading.Event() for _ in range(args.steps)]
        gathered = [threading.Event() for _ in range(args.steps)]
        step_ms, h2d_ms = [0.0] * args.steps, [0.0] * args.steps
        errors = []

        def worker(i):
            try:
                for k in range(i, args.steps, nfl):
                    if k - nfl >= 0:
                        gathered[k - nfl].wait()              # the results buffer of this context is free again
                    engs[i].run(plan, x_host, ress[i])
                    tm = engs[i].timings()
                    step_ms[k], h2d_ms[k] = tm["total"], tm["h2d"]
                    done[k].set()
            except BaseException as exc:                      # noqa: BLE001 - re-raised in the main thread
                errors.append(exc)
                for ev in done:
                    ev.set()

        w0 = time.perf_counter()
        threads = [threading.Thread(target=worker, args=(i,)) for i in range(nfl)]
        for th in threads:
            th.start()
        for k in range(args.steps):
            done[k].wait()
            if errors:
                break
            if pg is not None:
                import torch
                r_ = ress[k % nfl]
                for i, kk in enumerate(("mdccm", "mad", "vel", "baz", "sigma_tau")):
                    pg.local[i, :e_local].copy_(torch.from_numpy(getattr(r_, kk)[:e_local]), non_blocking=True)
                pg.gather()
                torch.cuda.synchronize()
            gathered[k].set()
        for ev in gathered:
            ev.set()
        for th in threads:
            th.join()
        if errors:
            raise errors[0]
        barrier()
        e_wall = (time.perf_counter() - w0) * 1e3
        e_ms, h_ms = sum(step_ms), sum(h2d_ms)
        if dist is not None:
            import torch
            tt = torch.tensor([e_wall, h_ms], dtype=torch.float64, device="cuda")
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            e_wall, h_ms = float(tt[0]), float(tt[1])
        e2e = {"value": e_total / (e_wall / args.steps * 1e-3), "unit": "estimates/s",
               "h2d_bytes_per_step": int(x_host.nbytes) * world, "d2h_bytes_per_step": int(res.nbytes()) * world,
               "ms_per_step_wall": e_wall / args.steps, "ms_record_latency_device": e_ms / args.steps,
               "records_in_flight": nfl,
               "h2d_gbs_per_rank": (x_host.nbytes / 1e9) / (h_ms / args.steps * 1e-3) if h_ms > 0 else None,
               "host_numa_node_rank0": numa,
               "note": (f"{nfl} records in flight, one context each: the H2D copy, the D2H copies and the host-side waits of "
                        "one record overlap the kernels of the others; one record alone spends ms_record_latency_device "
                        "in flight") if nfl > 1 else None}
        for e_ in engs[1:]:
            e_.close()

    check = None
    if world > 1:
        check = multi_gpu_check(j, eng, rank, world, dist, sos, method, args)
    return dict(j=j, plan=plan, world=world, rank=rank, value=value, ms_per_step=ms_per_step, wall_ms=wall_ms,
                stage={k: v / args.steps for k, v in stage.items()}, gather_ms=gather_ms / args.steps, clocks=clocks,
                launches=launches, e2e=e2e, fp64_peak=fp64_peak, fp32_peak=fp32_peak, xstats=xstats, e_total=e_total,
                e_local=e_local, shard=shard, g_npts=g_npts, dist=dist, keep=keep, prof=prof, eng=eng, check=check,
                method=method)


C5_ALPHAS = (1.0, 0.75)       # the OLS + LTS sweep of config 5


def run_batch(args):
    """c5: a batch of independent arrays, dealt to the ranks by array first (then by time segment when ranks
    outnumber arrays); every array is filtered and cross-correlated once and solved for every alpha of the sweep."""
    from narrow_band_least_squares_b200 import _capi
    from narrow_band_least_squares_b200.batch import deal_arrays
    from narrow_band_least_squares_b200.engine import Engine, Results, window_grid
    from narrow_band_least_squares_b200.helpers import band_edges, design_sos
    from narrow_band_least_squares_b200.sharding import Shard, filter_halo
    from narrow_band_least_squares_b200.synth import CONFIGS, blocked_record_cuda
    import torch

    world, rank, local_rank, dist = dist_setup()
    j = workload(args)
    c = CONFIGS[j["name"]]
    method = xcorr_method(args)
    fs, nchan, nb, narrays = j["fs"], j["nchans"], j["nbands"], j["narrays"]
    g_npts = int(round(j["hours_total"] * 3600.0 * fs))
    sos = np.stack([design_sos(j["filter_type"], *band_edges(j["freqlist"], b, "log"), fs, j["filter_order"],
                               j["filter_ripple"]) for b in range(nb)])
    grids = [window_grid(g_npts, fs, j["winlen_list"][b], j["winover"]) for b in range(nb)]
    win_len = [g[0] for g in grids]; win_hop = [g[1] for g in grids]; win_cnt = [g[2] for g in grids]
    dealt = deal_arrays(narrays, world)
    halo = filter_halo(sos) if narrays < world else 0
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(local_rank)
    eng = Engine(local_rank)
    if args.single_stage:
        eng.set_option(_capi.OPT_TWO_STAGE, 0)
    items = []
    for a, s, g in dealt[rank]:
        sh = Shard(g_npts, win_len, win_hop, win_cnt, s, g, halo)
        ld = (sh.npts + 15) // 16 * 16
        x, rij, _, _ = blocked_record_cuda(c["seed"] + 1000 * a, nchan, sh.seg_lo, sh.seg_hi, fs, c["aperture_km"], dev,
                                           ld=ld, flip_channel=3 if a % 2 else None)
        plan = sh.plan(nchan, fs, sos, True, win_len, win_hop, rij, alpha=C5_ALPHAS[0], max_lag=args.max_lag,
                       method=method)
        items.append(dict(a=a, ld=ld, x=_DevPtr(x), plan=plan, plans=[plan] + [plan.with_alpha(al) for al in C5_ALPHAS[1:]]))
    torch.cuda.synchronize()
    torch.cuda.empty_cache()
    e_local = sum(it["plan"].nwin for it in items) * len(C5_ALPHAS)
    e_total = narrays * int(sum(win_cnt)) * len(C5_ALPHAS)
    # outputs: one packed estimate block per (item, alpha) in a single tensor, so that the gather is one collective
    n_items_max = max(len(d) for d in dealt)
    e_max = max(int(sum(win_cnt)), 1)
    P = items[0]["plan"].npairs if items else nchan * (nchan - 1) // 2
    block = torch.zeros((n_items_max, len(C5_ALPHAS), 7, e_max), dtype=torch.float64, device=dev)
    allb = torch.empty((world,) + tuple(block.shape), dtype=torch.float64, device=dev) if world > 1 else None
    lag_t = torch.empty((e_max, P), dtype=torch.int32, device=dev)
    cmax_t = torch.empty((e_max, P), dtype=torch.float64, device=dev)
    w_t = torch.empty((e_max, P), dtype=torch.uint8, device=dev)

    def outputs(i, k):
        o = _capi.Outputs()
        base = block[i, k].data_ptr()
        o.mdccm, o.mad, o.vel, o.baz, o.sigma_tau, o.slowness = (base + q * 8 * e_max for q in range(6))
        o.lag, o.cmax, o.weights = lag_t.data_ptr(), cmax_t.data_ptr(), w_t.data_ptr()
        return o

    outs = [[outputs(i, k) for k in range(len(C5_ALPHAS))] for i in range(len(items))]

    def barrier():
        eng.synchronize()
        if dist is not None:
            torch.cuda.synchronize()
            dist.barrier()

    def step_resident():
        stage = {"filter": 0.0, "xcorr": 0.0, "solve": 0.0}
        eng.mark(0)
        for i, it in enumerate(items):
            eng.run_resident(it["plan"], it["x"], it["ld"], outs[i][0])
            for k in range(1, len(C5_ALPHAS)):
                eng.solve_resident(it["plans"][k], it["plan"].nwin, lag_t.data_ptr(), cmax_t.data_ptr(), outs[i][k])
            t = eng.timings()                      # waits for the run_device part of this array
            for kk in stage:
                stage[kk] += t[kk]
        eng.mark(1)
        ms = eng.elapsed(0, 1)
        stage["solve"] = ms - stage["filter"] - stage["xcorr"]      # both solves of every array
        mg = 0.0
        if allb is not None:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            dist.all_gather_into_tensor(allb.view(world * n_items_max, len(C5_ALPHAS), 7, e_max), block)
            e1.record()
            torch.cuda.synchronize()
            mg = e0.elapsed_time(e1)
        return ms, mg, stage

    fp64_peak = eng.fp64_peak_tflops(20000)
    fp32_peak = eng.fp32_peak_tflops(20000)
    sampler = ClockSampler(local_rank).start()
    for _ in range(args.warmup):
        step_resident()
    barrier()
    launches0 = eng.launch_count()
    eng.xcorr_stats(reset=True)
    sampler.mark()
    wall0 = time.perf_counter()
    dev_ms, gather_ms, stage = 0.0, 0.0, {"filter": 0.0, "xcorr": 0.0, "solve": 0.0}
    for _ in range(args.steps):
        ms, mg, st = step_resident()
        dev_ms += ms + mg
        gather_ms += mg
        for k in stage:
            stage[k] += st[k]
    barrier()
    wall_ms = (time.perf_counter() - wall0) * 1e3
    clocks = sampler.stop()
    launches = eng.launch_count() - launches0
    xstats = eng.xcorr_stats()
    xstats["pair_windows"] = int(sum(it["plan"].nwin for it in items) * P * args.steps)
    if dist is not None:
        tt = torch.tensor([dev_ms, wall_ms, gather_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dev_ms, wall_ms, gather_ms = float(tt[0]), float(tt[1]), float(tt[2])
    ms_per_step = dev_ms / args.steps
    value = e_total / (ms_per_step * 1e-3)

    prof = None
    if not args.no_profile and items:
        eng.set_option(_capi.OPT_PROFILE, 1)
        eng.profile(reset=True)
        step_resident()
        prof = eng.profile(reset=True)
        eng.set_option(_capi.OPT_PROFILE, 0)

    # e2e: host buffers through nbls_run_host + nbls_resolve_host, `nfl` arrays in flight on their own contexts.  The
    # arrays of a rank cycle through at most 4 distinct pinned host records (64 x 553 MB would pin 35 GB): every array
    # still copies its input from pinned host memory and its results back inside the timed region.
    e2e = None
    if not args.no_e2e and items:
        nfl = max(1, min(args.e2e_in_flight, len(items)))
        engs = [eng] + [Engine(local_rank) for _ in range(nfl - 1)]
        if args.single_stage:
            for e_ in engs[1:]:
                e_.set_option(_capi.OPT_TWO_STAGE, 0)
        nh = min(4, len(items))
        hosts = []
        for i in range(nh):
            h = _capi.pinned_empty((nchan, items[i]["plan"].npts))
            torch.from_numpy(h).copy_(items[i]["x"].tensor[:, :items[i]["plan"].npts])
            hosts.append(h)
        ress = [[Results(p, pinned=True, want_cmax=False) for p in items[0]["plans"]] for _ in range(nfl)]
        d2h = sum(r.nbytes() for r in ress[0]) - ress[0][1].lag.nbytes

        def run_item(e_, slot, i):
            it = items[i]
            e_.run(it["plan"], hosts[i % nh] if it["plan"].npts == hosts[i % nh].shape[1] else hosts[0], ress[slot][0])
            for k in range(1, len(C5_ALPHAS)):
                e_.resolve(it["plans"][k], ress[slot][k])

        for s_, e_ in enumerate(engs):
            run_item(e_, s_, 0)
        barrier()
        errors = []

        def worker(s_):
            try:
                for _ in range(args.steps):
                    for i in range(s_, len(items), nfl):
                        run_item(engs[s_], s_, i)
            except BaseException as exc:      # noqa: BLE001
                errors.append(exc)

        w0 = time.perf_counter()
        threads = [threading.Thread(target=worker, args=(s_,)) for s_ in range(nfl)]
        for th in threads:
            th.start()
        for th in threads:
            th.join()
        if errors:
            raise errors[0]
        if allb is not None:
            dist.all_gather_into_tensor(allb.view(world * n_items_max, len(C5_ALPHAS), 7, e_max), block)
        barrier()
        e_wall = (time.perf_counter() - w0) * 1e3
        if dist is not None:
            tt = torch.tensor([e_wall], dtype=torch.float64, device="cuda")
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            e_wall = float(tt[0])
        e2e = {"value": e_total / (e_wall / args.steps * 1e-3), "unit": "estimates/s",
               "h2d_bytes_per_step": int(sum(8 * nchan * it["plan"].npts for it in items)) * world,
               "d2h_bytes_per_step": int(d2h) * len(items) * world, "ms_per_step_wall": e_wall / args.steps,
               "records_in_flight": nfl,
               "note": f"{nfl} arrays in flight, one context each; arrays cycle through {nh} pinned host records"}
        for e_ in engs[1:]:
            e_.close()
    plan0 = items[0]["plan"] if items else None
    return dict(j=j, plan=plan0, world=world, rank=rank, value=value, ms_per_step=ms_per_step, wall_ms=wall_ms,
                stage={k: v / args.steps for k, v in stage.items()}, gather_ms=gather_ms / args.steps, clocks=clocks,
                launches=launches, e2e=e2e, fp64_peak=fp64_peak, fp32_peak=fp32_peak, xstats=xstats, e_total=e_total,
                e_local=e_local, g_npts=g_npts, dist=dist, keep=(block, allb, lag_t, cmax_t, w_t, items), prof=prof,
                eng=eng, check=None, method=method, n_items=len(items))


GROUP_BOUND = {  # kernel group -> (bound, unit of `work`)
    "filter_fwd_phase1": "hbm", "filter_fwd_phase3": "hbm", "filter_bwd_phase1": "hbm", "filter_bwd_phase3": "hbm",
    "filter_carry": "latency", "xf_spectra": "fp32", "xf_pair": "fp32", "xf_verify": "fp64", "xf_verify_items": "fp64",
    "xcorr_direct": "fp64", "stats_ols": "latency", "lts": "issue"}


def kernel_groups_block(prof, peaks, fp32_peak, fp64_peak, traffic):
    """Per kernel group of the profiled step: own CUDA-event time, executed algorithmic work, fraction of ITS bound."""
    total = sum(v["ms"] for v in prof.values()) or 1.0
    out = {}
    for name, v in prof.items():
        if v["launches"] == 0:
            continue
        bound = GROUP_BOUND.get(name, "latency")
        g = {"ms": v["ms"], "launches": v["launches"], "share": v["ms"] / total, "bound": bound}
        if v["work"] > 0 and v["ms"] > 0:
            if bound == "hbm":
                g.update(achieved=v["work"] / (v["ms"] * 1e-3) / 1e9, unit="GB/s", peak=peaks["hbm_gbs"],
                         algorithmic_bytes=v["work"])
            elif bound in ("fp32", "fp64"):
                g.update(achieved=v["work"] / (v["ms"] * 1e-3) / 1e12, unit="TFLOP/s",
                         peak=fp32_peak if bound == "fp32" else fp64_peak, executed_flop=v["work"])
            if "achieved" in g:
                g["frac"] = g["achieved"] / g["peak"]
        if name in traffic:
            g["traffic"] = traffic[name]
        out[name] = g
    return out


def config_block(j, world, args, extra=None):
    per = j["hours_total"] if args.scaling == "strong" or world == 1 else j["hours_total"]
    if j["name"] == "c5":
        shard = (f"{j['narrays']} independent arrays dealt to {world} GPU(s) by array"
                 + (" then by time segment" if j["narrays"] < world else ""))
        wl = (f"c5: batch of {j['narrays']} arrays x {j['nchans']} elements, {j['hours_total']:g} h at {j['fs']:g} sps each, "
              f"{j['nbands']} log bands 0.5-5 Hz, OLS + LTS(alpha={C5_ALPHAS[1]:g}) sweep on one cross-correlation pass, "
              f"{j['winlen_list'][0]:g} s windows, 50% overlap; an estimate = one (band-window, alpha)")
    else:
        if world == 1:
            shard = "1 GPU"
        elif args.scaling == "weak":
            shard = f"time segments with filter halos over {world} GPUs, record = {world} x {per:g} h"
        else:
            shard = f"ONE record of {per:g} h split into {world} time segments with filter halos"
        wl = (f"{j['name']}: {j['nchans']}-element array, {per:g} h at {j['fs']:g} sps"
              + (" per GPU" if args.scaling == "weak" and world > 1 else "") +
              f", {j['nbands']} log bands 0.5-5 Hz, alpha={j['alpha']:g}, windows {j['winlen_list'][0]:g}->"
              f"{j['winlen_list'][-1]:g} s, 50% overlap, butter order {j['filter_order']} zero-phase")
    c = {"workload": wl, "sharding": shard, "l2": "inputs larger than L2 (record and filtered bands are GB-sized)",
         "xcorr_method": args.method + (" (single-stage)" if args.single_stage else ""), "max_lag": args.max_lag,
         "record": "incoherent noise (plane wave 30 dB down)" if args.noise_only else "synthetic plane wave + noise, every 10th minute noise-only",
         "e2e": (f"up to {max(1, min(args.e2e_in_flight, args.steps))} record(s) in flight (see e2e.records_in_flight), one "
                 "context each; every step copies its input from pinned host memory and its results back inside the "
                 "timed region")}
    if extra:
        c.update(extra)
    return c


def reference_arm(args, ncores):
    """--impl reference: the reference's own CPU driver on the host cores, bounded slice per step."""
    j = workload(args)
    alphas = C5_ALPHAS if j["name"] == "c5" else (j["alpha"],)
    npts = int(round(args.cpu_sample_min * 60.0 * j["fs"]))
    rec = make_record(j, npts)
    kind = "reference"

    def one_step():
        nonlocal kind
        est, secs = 0, 0.0
        for al in alphas:
            jj = dict(j, alpha=al)
            r = cpu_reference_run(jj, rec)
            if r is None:
                kind = "port"
                out, dt = cpu_port_run(jj, rec, ncores)
                est += int(sum(out["num_compute"]))
            else:
                out, dt = r
                est += int(sum(out[6]))
            secs += dt
        return est, secs

    cold = None
    for _ in range(max(0, args.warmup)):
        e, s_ = one_step()
        cold = cold if cold is not None else e / s_
        if s_ * (args.warmup + args.steps) > 240:     # keep the whole arm within a few minutes
            break
    rates, est, secs = [], 0, 0.0
    for _ in range(args.steps):
        est, dt = one_step()
        rates.append(est / dt); secs += dt
        if secs > 200:
            break
    v = float(np.mean(rates))
    sample = (f"{args.cpu_sample_min:g}-minute slice of the workload, all {j['nbands']} bands, {est} estimates per step; "
              + ("the reference's own narrow_band_least_squares_parallel (joblib n_jobs=-1), unmodified, via oracle/refstubs"
                 if kind == "reference" else "oracle port (staged reference absent)")
              + (f"; first (cold) call {cold:.1f} est/s" if cold else ""))
    print(json.dumps({"impl": "reference", "metric": "NBLS band-window estimates/sec", "value": v,
                      "unit": "estimates/s", "n_gpus": args.gpus, "steps": len(rates), "warmup": args.warmup,
                      "ms_per_step": secs / len(rates) * 1e3, "higher_is_better": True, "scaling": args.scaling,
                      "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                      "config": config_block(j, 1, args),
                      "cpu_baseline": {"value": v, "unit": "estimates/s", "cores": ncores, "kind": kind,
                                       "cold_value": cold, "sample": sample},
                      "e2e": {"value": v, "unit": "estimates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    ncores = os.cpu_count() or 1
    if args.impl == "reference":
        if rank == 0:
            reference_arm(args, ncores)
        return
    r = run_batch(args) if args.workload == "c5" else run_single(args)
    if r["rank"] != 0:
        if r["dist"] is not None:
            r["dist"].destroy_process_group()
        return
    j, plan, world, eng = r["j"], r["plan"], r["world"], r["eng"]
    peaks, peak_src = measured_peaks()
    from narrow_band_least_squares_b200 import _capi
    lib = _capi.load()
    B, C, P = plan.nbands, plan.nchan, plan.npairs
    n_local_arrays = r.get("n_items", 1)
    T = plan.npts
    # single-stage-equivalent work of this rank's step: what the path would execute without the two-stage search
    # (the figure BENCH_r01 called `roofline`), and the reference algorithm's own count (np.correlate: P N^2 FMA)
    flop_single, flop_ref = 0.0, 0.0
    for b in range(B):
        W, N = int(plan.win_count[b]), int(plan.win_len[b])
        L = int(lib.nbls_xcorr_fft_length(N, args.max_lag)) if (args.method != "direct" and N >= 128) else 0
        if L:
            flop_single += W * (((C + 1) // 2 + (P + 1) // 2) * 5.0 * L * np.log2(L) + ((P + 1) // 2) * L * 16.0)
        flop_ref += W * P * 2.0 * float(N) ** 2
    flop_single *= n_local_arrays
    flop_ref *= n_local_arrays
    x_ms, f_ms = r["stage"]["xcorr"], r["stage"]["filter"]
    key = f"{j['name']}_{j['hours_total']:g}h_{args.method}" + ("_1stage" if args.single_stage else "")
    traffic = ncu_traffic(key) if world == 1 else {}
    groups = kernel_groups_block(r["prof"], peaks, r["fp32_peak"], r["fp64_peak"], traffic) if r["prof"] else {}
    roof = {"kernel": None}
    if groups:
        dom = max(groups, key=lambda k: groups[k]["ms"])
        g = groups[dom]
        xg = [k for k in groups if k.startswith("xf_") or k == "xcorr_direct"]
        x_flop32 = sum(groups[k].get("executed_flop", 0.0) for k in xg if groups[k]["bound"] == "fp32")
        x_flop64 = sum(groups[k].get("executed_flop", 0.0) for k in xg if groups[k]["bound"] == "fp64")
        if r.get("xstats") and r["xstats"].get("verified"):
            # lazy verification sizes its item lists on the device, so the library cannot count that work on the host:
            # candidates evaluated exactly per step (xcorr_stats) x 2 N flop at the window-weighted mean window length
            wsum = float(sum(int(plan.win_count[b]) for b in range(B)))
            n_mean = sum(int(plan.win_count[b]) * float(plan.win_len[b]) for b in range(B)) / max(wsum, 1.0)
            x_flop64 = max(x_flop64, r["xstats"]["verified"] / max(args.steps, 1) * 2.0 * n_mean)
        roof = {"kernel": dom, "bound": g["bound"], "achieved": g.get("achieved"), "peak": g.get("peak"),
                "unit": g.get("unit"), "frac": g.get("frac"), "traffic": g.get("traffic"),
                "ms_per_step": g["ms"], "share_of_step": g["share"], "launches_per_step": g["launches"],
                "how": ("executed algorithmic flop of the group's launches in one profiled step (per band: transform "
                        "length actually run, windows and pair-windows actually redone) / the group's own CUDA-event "
                        "time on libnbls' stream"),
                "peak_source": ("FP32 FMA microbenchmark run in this process, 2 flop per FMA.  The kernel is bound by "
                                "the FP32 issue pipe: HBM traffic is a few % of the copy peak and tcgen05 has no FP32/"
                                "FP64 mode (DESIGN.md section 4 holds the measured tensor-pipe experiment), so neither "
                                "MEASURED_PEAKS.json entry is its roofline.  FFT butterflies are mostly unfused "
                                "adds: 0.5 is the ceiling of this fraction") if g["bound"] == "fp32" else peak_src,
                "xcorr_stage": {"executed_fp32_flop": x_flop32, "executed_fp64_flop": x_flop64, "ms_timed": x_ms,
                                "achieved_fp32_tflops": x_flop32 / (x_ms * 1e-3) / 1e12 if x_ms > 0 else None,
                                "frac_of_fp32_peak": x_flop32 / (x_ms * 1e-3) / 1e12 / r["fp32_peak"] if x_ms > 0 else None},
                "algorithmic_speedup": {
                    "single_stage_flop": flop_single, "executed_over_single_stage": (x_flop32 / flop_single) if flop_single else None,
                    "reference_algorithm_flop": flop_ref,
                    "reference_equivalent_tflops": flop_ref / (x_ms * 1e-3) / 1e12 if x_ms > 0 else None,
                    "note": "the reference computes every lag by np.correlate (P N^2 multiply-adds per window, SURVEY 8d); "
                            "the FFT screen + exact candidates returns the same integer lags with far fewer operations - an "
                            "algorithmic gain, not a roofline fraction"},
                "fp64_peak": r["fp64_peak"], "fp32_peak": r["fp32_peak"]}
    filt_bytes = 8.0 * C * T * (1 + 3 * B) * n_local_arrays
    fgb = filt_bytes / (f_ms * 1e-3) / 1e9 if f_ms > 0 else None
    ftraffic = sum(v for k, v in traffic.items() if k.startswith("filter_")) or None
    line = {"metric": "NBLS band-window estimates/sec", "value": r["value"], "unit": "estimates/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"],
            "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_block(j, world, args, {"estimates_per_step": r["e_total"]}),
            "clocks": r["clocks"], "gpu_launches": int(r["launches"]), "e2e": r["e2e"],
            "stage_ms": r["stage"], "gather_ms": r["gather_ms"], "wall_ms_per_step": r["wall_ms"] / args.steps,
            "roofline": roof, "kernel_groups": groups,
            "kernel_groups_note": ("one extra step with NBLS_OPT_PROFILE: a CUDA-event pair around every kernel group, the "
                                   "filter's two band groups run one after the other there (in the timed steps they overlap "
                                   "on two streams, so stage_ms.filter is smaller than the sum of its groups)"),
            "roofline_filter": {"kernel": "biquad_state_shared_kernel + biquad_chunk_kernel x3 + carry_kernel x2",
                                "bound": "hbm", "achieved": fgb, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                                "frac": (fgb / peaks["hbm_gbs"]) if fgb else None, "peak_source": peak_src,
                                "algorithmic_bytes": filt_bytes, "traffic": ftraffic},
            "traffic_source": ("profiles/ncu_traffic.json[" + key + "] (scripts/ncu_traffic.sh)") if traffic else None,
            "xcorr_stats": r["xstats"]}
    if r["check"] is not None:
        line["multi_gpu_check"] = r["check"]
    if not args.no_cpu and world == 1:
        alphas = C5_ALPHAS if j["name"] == "c5" else (j["alpha"],)
        bases, par = [], {}
        for al in alphas:
            jj = dict(j, alpha=al)
            base, rec, ref_out, port_out = cpu_legs(jj, args.cpu_sample_min, ncores)
            bases.append(base)
            pb = parity_block(eng, jj, rec, ref_out, port_out, args, r["method"]) if len(alphas) == 1 else \
                parity_block_alpha(eng, jj, rec, ref_out, port_out, args, r["method"], al)
            par[f"alpha={al:g}"] = pb
        if len(bases) == 1:
            line["cpu_baseline"], line["parity"] = bases[0], par[f"alpha={alphas[0]:g}"]
        else:       # sweep: estimates and seconds of both passes add up
            tot = sum(1.0 / b["value"] for b in bases)
            line["cpu_baseline"] = {"value": len(bases) / tot / 1.0, "unit": "estimates/s", "cores": ncores,
                                    "kind": bases[0]["kind"],
                                    "sample": " | ".join(f"alpha={a:g}: {b['sample']}" for a, b in zip(alphas, bases))}
            line["parity"] = par
    print(json.dumps(line))
    if r["dist"] is not None:
        r["dist"].destroy_process_group()


def parity_block_alpha(eng, jj, rec, ref_out, port_out, args, method, alpha):
    """parity_block for one alpha of a sweep workload (the config's own alpha is the first of the sweep)."""
    import narrow_band_least_squares_b200.configs as cfg
    orig = cfg.config_job

    def patched(name):
        d = orig(name)
        d["alpha"] = alpha
        return d
    cfg.config_job = patched
    try:
        return parity_block(eng, jj, rec, ref_out, port_out, args, method)
    finally:
        cfg.config_job = orig


if __name__ == "__main__":
    main()
